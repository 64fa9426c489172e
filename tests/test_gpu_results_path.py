"""Results written straight to page-locked host memory must equal the copied ones:
  * the keypoint store (compacted by libsift3D.so from the in-place candidate view,
    sift3d_cuda_candidates_view) against the reference's compaction rule (sift.c:1148-1166,
    strength slot quirk of copy_Keypoint sift.c:372-384) applied to the records that
    sift3d_cuda_fetch_candidates copies out;
  * descriptors written by the kernel into the pinned store against descriptors written to a
    device buffer and copied into pageable memory (sift3d_cuda_describe with a numpy array)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_store_equals_compaction_of_fetched_candidates_and_direct_descriptors():
    import sift3d_b200 as s3
    from sift3d_b200 import api, synth

    vol = synth.synth_volume(112, 96, 80, seed=21, flags=3)
    det = s3.Detector(0)
    kp = det.detect(s3.Image.from_array(vol))
    rec = kp.records()
    cand = det.candidates()  # copy path
    valid = cand[cand["valid"] != 0]
    assert len(cand) > 50 and len(valid) == len(rec) > 10
    for f in ("o", "s", "x", "y", "z", "sd"):
        assert np.array_equal(rec[f], valid[f]), f
    assert np.array_equal(rec["R"].view(np.uint32), valid["R"].reshape(len(valid), 9).view(np.uint32))
    # survivor j carries the strength of candidate j
    assert np.array_equal(rec["strength"].view(np.uint32), cand["strength"][:len(rec)].view(np.uint32))

    direct = det.extract(kp).to_matrix()[:, 3:]  # kernel -> pinned store
    L = api.libcuda()
    keys = np.ascontiguousarray(valid)
    out = np.full((len(keys), 768), np.nan, dtype=np.float32)  # pageable: device buffer + copy
    L.sift3d_cuda_describe.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int]
    L.sift3d_cuda_describe.restype = C.c_int
    assert L.sift3d_cuda_describe(det.ctx, keys.ctypes.data_as(C.c_void_p), len(keys),
                                  out.ctypes.data_as(C.c_void_p), 0) == 0
    assert np.array_equal(direct.view(np.uint32), out.view(np.uint32))
    det.close()
