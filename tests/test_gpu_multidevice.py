"""Detectors on different devices of ONE process (sift3d_detector_set_device / Detector(dev)):
kernel attributes and __constant__ tables are per device, so the second device must be set up
too and give the bits of the first. Needs two GPUs; skipped on a one-GPU box."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _bits(a):
    return np.ascontiguousarray(a).view(np.uint8)


def test_second_device_in_same_process_matches_first():
    import sift3d_b200 as s3
    from sift3d_b200 import synth

    if s3.cuda_device_count() < 2:
        pytest.skip("needs two CUDA devices")
    vol = synth.synth_volume(96, 80, 72, seed=11, flags=3)
    out = []
    for dev in (0, 1):
        det = s3.Detector(dev)
        kp = det.detect(s3.Image.from_array(vol))
        cands = det.candidates()
        desc = det.extract(kp).to_matrix()
        out.append((cands, kp.records(), desc))
        det.close()
    (c0, k0, d0), (c1, k1, d1) = out
    assert len(c0) > 20 and len(k0) > 5
    assert len(c0) == len(c1) and len(k0) == len(k1)
    for f in c0.dtype.names:
        assert np.array_equal(_bits(c0[f]), _bits(c1[f])), f
    assert np.array_equal(_bits(d0), _bits(d1))
