"""sift3d_image_prefetch: the asynchronous upload consumed by the next detect must give the
result of the synchronous path, be single-use, and survive a changed volume shape."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _cands(det, im):
    det.detect(im)
    return det.candidates()


def _same(a, b):
    """Field-wise bit equality (the records carry padding bytes)."""
    return len(a) == len(b) and all(
        np.array_equal(np.ascontiguousarray(a[f]).view(np.uint8), np.ascontiguousarray(b[f]).view(np.uint8))
        for f in a.dtype.names)


def test_prefetched_detect_equals_host_detect():
    import sift3d_b200 as s3
    from sift3d_b200 import synth

    det = s3.Detector(0)
    vol = synth.synth_volume(96, 80, 64, seed=5, flags=3)
    im = s3.Image.from_array(vol)
    want = _cands(det, im)
    im.prefetch(0)
    got = _cands(det, im)
    assert len(want) > 20
    assert _same(got, want)

    # single use: without a new prefetch the host buffer is read again
    im.array[...] = synth.synth_volume(96, 80, 64, seed=6, flags=3)
    want2 = _cands(det, im)
    assert not _same(want2, want)
    im.prefetch(0)
    im.prefetch_wait()
    assert _same(_cands(det, im), want2)


def test_prefetch_pipeline_over_several_images():
    import sift3d_b200 as s3
    from sift3d_b200 import synth

    det = s3.Detector(0)
    shapes = [(64, 64, 64), (96, 64, 80), (64, 64, 64)]
    ims = [s3.Image.from_array(synth.synth_volume(*sh, seed=10 + i, flags=3)) for i, sh in enumerate(shapes)]
    want = [_cands(det, im).copy() for im in ims]
    ims[0].prefetch(0)
    for i, im in enumerate(ims):
        if i + 1 < len(ims):
            ims[i + 1].prefetch(0)
        got = _cands(det, im)
        assert _same(got, want[i])
