"""sift3d_detect_describe_sharded (C host + libsift3d_cuda multi.cu: z-slabs of one volume, one
host thread per slab, halo exchange by in-stream peer copies) must give the one-device result
bit for bit: keypoints (incl. the strength-slot quirk) and descriptors. On a one-GPU box all
slabs share device 0 (the peer copies degenerate to device-local ones); with more GPUs every
slab gets a device of its own. Blur arithmetic: exact (the slab kernels see other tile and
chunk shapes than the whole-volume ones; only the 0-ULP mode is shape independent)."""
import numpy as np
import pytest

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(600)]


def _single(vol, units, params, exact=True):
    import sift3d_b200 as s3
    det = s3.Detector(0)
    det.set_exact(exact)
    for k, v in (params or {}).items():
        assert getattr(det, f"set_{k}")(v) == 0
    kp = det.detect(s3.Image.from_array(vol, units))
    keys = kp.records()
    desc = det.extract(kp).to_matrix() if len(kp) else np.zeros((0, 771), np.float32)
    det.close()
    return keys, desc


def _sharded(vol, units, params, world, octaves):
    import sift3d_b200 as s3
    ndev = s3.cuda_device_count()
    det = s3.Detector(0)
    for k, v in (params or {}).items():
        assert getattr(det, f"set_{k}")(v) == 0
    devices = [r % ndev for r in range(world)]
    im = s3.Image.from_array(vol, units)
    out = None
    for _ in range(2):  # second run: the cached multi-device object is reused
        kp, ds = det.detect_describe_sharded(im, devices, octaves)
        cur = (kp.records(), ds.to_matrix() if len(kp) else np.zeros((0, 771), np.float32))
        if out is not None:
            assert np.array_equal(out[1].view(np.uint32), cur[1].view(np.uint32))
        out = cur
    tm = det.sharded_timings()
    det.close()
    return out[0], out[1], tm


@pytest.fixture(autouse=True)
def _exact_mode(monkeypatch, request):
    # contexts made by the library read the variable at creation
    if "fast_mode" in request.keywords:
        monkeypatch.delenv("SIFT3D_CUDA_EXACT", raising=False)
    else:
        monkeypatch.setenv("SIFT3D_CUDA_EXACT", "1")


@pytest.mark.parametrize("world,dims,octaves,units,params", [
    (2, (96, 96, 320), 2, (1.0, 1.0, 1.0), None),      # two sharded octaves + unsharded tail
    (2, (128, 64, 250), 1, (1.0, 1.0, 1.0), None),     # nz not a multiple of the alignment
    (3, (64, 64, 480), 2, (1.0, 1.0, 1.0), None),      # an interior slab with two cut faces
    (4, (64, 64, 640), 2, (1.0, 1.0, 1.0), None),
    (8, (64, 64, 512), 3, (1.0, 1.0, 1.0), None),      # halos that reach the volume's faces
    (8, (96, 64, 512), 4, (1.0, 1.0, 1.0), None),      # every octave sharded (no unsharded tail)
    (2, (64, 64, 384), 2, (2.0, 2.0, 2.0), {"peak_thresh": 0.05, "corner_thresh": 0.3, "num_kp_levels": 4}),
    (2, (64, 64, 320), 1, (0.9, 0.9, 1.4), None),      # non-dyadic tap spacing
])
def test_sharded_equals_single_device(world, dims, octaves, units, params):
    from sift3d_b200.synth import synth_volume
    nx, ny, nz = dims
    vol = synth_volume(nx, ny, nz, num_blobs=nx * ny * nz // 4096, seed=77 + nz, flags=3)
    keys, desc = _single(vol, units, params)
    got_keys, got_desc, tm = _sharded(vol, units, params, world, octaves)
    assert len(keys) > 10
    assert len(got_keys) == len(keys)
    for f in keys.dtype.names:
        assert np.array_equal(got_keys[f], keys[f]), f
    assert np.array_equal(got_desc.view(np.uint32), desc.view(np.uint32))
    assert tm["exchanged_mib"] > 0


@pytest.mark.fast_mode
@pytest.mark.parametrize("world,dims,octaves,units", [
    (2, (96, 96, 320), 2, (1.0, 1.0, 1.0)),      # fused kernels on the slabs, banded kernels below
    (4, (64, 64, 640), 2, (1.0, 1.0, 1.0)),
    (2, (64, 64, 320), 1, (0.9, 0.9, 1.4)),      # banded kernels with a z origin (cut faces)
])
def test_sharded_fast_mode_matches_single_device(world, dims, octaves, units):
    """The default (fast) blur arithmetic on the slabs: one FFMA per tap in the fused kernels, the
    banded kernels with band tables built for a slab's z origin. The slab kernels see other tile
    and chunk shapes than the whole-volume ones, so the comparison is the fast mode's own:
    same keypoints (>= 99 % matched by position and scale) and descriptors of cosine >= 0.999."""
    import sys, os
    sys.path.insert(0, os.path.dirname(__file__))
    import parity
    from sift3d_b200.synth import synth_volume
    nx, ny, nz = dims
    vol = synth_volume(nx, ny, nz, num_blobs=nx * ny * nz // 4096, seed=77 + nz, flags=3)
    keys, desc = _single(vol, units, None, exact=False)
    got_keys, got_desc, tm = _sharded(vol, units, None, world, octaves)
    assert len(keys) > 10
    pairs = parity.match_keypoints(keys, got_keys)
    assert len(pairs) >= 0.99 * len(keys) and abs(len(got_keys) - len(keys)) <= 0.01 * len(keys) + 1
    iw = [i for i, _ in pairs]
    ig = [j for _, j in pairs]
    cos = parity.cosine_rows(got_desc[ig, 3:], desc[iw, 3:])
    assert cos.min() >= parity.DESC_COS_MIN, cos.min()
    assert tm["exchanged_mib"] > 0
