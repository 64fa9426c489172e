"""kpSift3D (cli/kpSift3D.c, the reference's command-line program on the drop-in API) and
sift3d_read_image on a real NIfTI file: the CLI must write exactly the CSVs that the same call
sequence gives through the library (kpSift3D.c:96-146: read, detect, keep the 100 strongest,
save keypoints, extract, save descriptors), and its keypoints must be the oracle's."""
import ctypes as C
import gzip
import os
import subprocess

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "sift3d_b200", "bin", "kpSift3D")


def _write_nii(path, vol, units):
    from sift3d_b200 import api

    L = api.lib()
    L.sift3d_nifti_write_raw.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_double),
                                         C.POINTER(C.c_float)]
    nz, ny, nx = vol.shape
    assert L.sift3d_nifti_write_raw(os.fsencode(path), (C.c_int * 4)(nx, ny, nz, 1),
                                    (C.c_double * 3)(*units),
                                    vol.ctypes.data_as(C.POINTER(C.c_float))) == 0


def test_read_image_gives_the_volume_and_units(tmp_path):
    import sift3d_b200 as s3
    from sift3d_b200 import api, synth

    vol = synth.synth_volume(40, 36, 32, seed=3, flags=3)
    path = tmp_path / "v.nii.gz"
    _write_nii(path, vol, (0.75, 0.75, 1.5))
    L = api.lib()
    L.sift3d_read_image.restype = C.c_void_p
    h = L.sift3d_read_image(os.fsencode(path))
    assert h
    dims = [C.c_int() for _ in range(4)]
    L.sift3d_image_dimensions(h, *[C.byref(d) for d in dims])
    assert [d.value for d in dims] == [40, 36, 32, 1]
    u = [C.c_double() for _ in range(3)]
    L.sift3d_image_get_units(h, *[C.byref(x) for x in u])
    assert [x.value for x in u] == [0.75, 0.75, 1.5]
    data = np.ctypeslib.as_array(L.sift3d_image_data(h), shape=vol.shape)
    assert data.tobytes() == vol.tobytes()
    L.sift3d_free_image.argtypes = [C.c_void_p]
    L.sift3d_free_image(h)
    assert not L.sift3d_read_image(os.fsencode(tmp_path / "v.txt"))
    assert not L.sift3d_read_image(os.fsencode(tmp_path / "absent.nii"))


def test_cli_matches_library_sequence_and_oracle(tmp_path):
    import sift3d_b200 as s3
    from sift3d_b200 import synth

    assert os.path.exists(CLI), "kpSift3D was not built (make)"
    vol, _ = synth.workload("C1_blobs_128")   # config 1: the CLI path
    nii = tmp_path / "c1.nii"
    _write_nii(nii, vol, (1.0, 1.0, 1.0))
    keys, desc = tmp_path / "keys.csv", tmp_path / "desc.csv.gz"
    r = subprocess.run([CLI, "--keys", str(keys), "--desc", str(desc), str(nii)],
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr

    # the same sequence through the library
    det = s3.Detector(0)
    im = s3.Image.from_array(vol)
    kp = det.detect(im)
    kp.sort_by_strength(100)
    k2, d2 = tmp_path / "keys2.csv", tmp_path / "desc2.csv.gz"
    assert kp.save(k2) == 0
    assert det.extract(kp).save(d2) == 0
    assert keys.read_bytes() == k2.read_bytes()
    assert gzip.open(desc).read() == gzip.open(d2).read()
    rows = np.loadtxt(keys, delimiter=",", ndmin=2)
    assert rows.shape == (100, 15)          # strength, x, y, z, o, sd, R (sift.c:1746-1788)

    # and against the reference's own run of that sequence
    want = parity.run_reference(vol, limit=100)
    got = parity.run_product(vol, limit=100)
    pairs = parity.match_keypoints(got["keys"], want["keys"])
    assert len(pairs) >= 0.99 * len(want["keys"])
    # the CSV holds octave coordinates, the matrix base-octave ones (sift.c:1651-1668)
    np.testing.assert_allclose(rows[:, 1:4] * np.exp2(rows[:, 4:5]), got["kp_matrix"], atol=1e-6)

    # argument errors of the reference program (kpSift3D.c:80-94)
    assert subprocess.run([CLI, str(nii)], capture_output=True).returncode == 1
    assert subprocess.run([CLI, "--keys", str(keys)], capture_output=True).returncode == 1
    assert subprocess.run([CLI, "--keys", str(keys), str(tmp_path / "none.nii")],
                          capture_output=True).returncode == 1
    assert subprocess.run([CLI, "--help"], capture_output=True).returncode == 0
