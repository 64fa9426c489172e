"""sift3d_detect_describe_batch (C-level batch entry, one host thread per detector) and the
upstream-cased aliases SIFT3D_detect_keypoints / SIFT3D_extract_descriptors: every volume's
result must equal the one-volume calls bit for bit. With one GPU the detectors share device 0
(two contexts, two host threads); with more, each detector gets a device of its own."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _single(vol, units):
    import sift3d_b200 as s3
    det = s3.Detector(0)
    kp = det.detect(s3.Image.from_array(vol, units))
    out = (kp.records(), det.extract(kp).to_matrix() if len(kp) else None)
    det.close()
    return out


def test_batch_equals_single_volume_calls():
    import sift3d_b200 as s3
    from sift3d_b200.synth import synth_volume
    ndev = s3.cuda_device_count()
    dets = [s3.Detector(d % ndev) for d in range(2 if ndev < 2 else min(ndev, 4))]
    vols = [synth_volume(64, 56, 48, num_blobs=40, seed=100 + i, flags=3) for i in range(5)]
    vols.append(np.zeros((48, 56, 64), np.float32))  # no keypoints: empty stores, no failure
    units = (1.0, 1.0, 1.0)
    images = [s3.Image.from_array(v, units) for v in vols]
    kps, dss = s3.detect_describe_batch(dets, images)
    for i, v in enumerate(vols):
        keys, desc = _single(v, units)
        got = kps[i].records()
        assert len(got) == len(keys)
        for f in keys.dtype.names:
            assert np.array_equal(got[f], keys[f]), (i, f)
        if desc is None:
            assert len(dss[i]) == 0
        else:
            assert np.array_equal(dss[i].to_matrix().view(np.uint32), desc.view(np.uint32)), i
    assert len(kps[-1]) == 0
    # detection only
    kps2, _ = s3.detect_describe_batch(dets, images[:3], describe=False)
    assert [len(k) for k in kps2] == [len(k) for k in kps[:3]]
    for d in dets:
        d.close()


def test_upstream_cased_aliases():
    import sift3d_b200 as s3
    from sift3d_b200.synth import synth_volume
    L = s3.lib()
    vol = synth_volume(64, 56, 48, num_blobs=40, seed=7)
    keys, desc = _single(vol, (1.0, 1.0, 1.0))
    det = s3.Detector(0)
    im = s3.Image.from_array(vol)
    kp, ds = s3.KeypointStore(), s3.DescriptorStore()
    assert L.SIFT3D_detect_keypoints(det.handle, im.handle, kp.handle) == 0
    assert L.SIFT3D_extract_descriptors(det.handle, kp.handle, ds.handle) == 0
    assert np.array_equal(kp.records()["R"], keys["R"])
    assert np.array_equal(ds.to_matrix().view(np.uint32), desc.view(np.uint32))
    det.close()


def test_four_pipelines_on_one_device_resident_inputs_and_reused_stores():
    """bench.py's pattern: P detectors on ONE device, images resident in HBM (the workers'
    sift3d_image_prefetch is then a no-op), every pipeline filling its own store again and again:
    the store of a pipeline ends up with its LAST volume, equal to the one-volume calls."""
    import sift3d_b200 as s3
    from sift3d_b200.synth import synth_volume
    P, rounds = 4, 3
    units = (1.0, 1.0, 1.0)
    vols = [synth_volume(64, 56, 48, num_blobs=40, seed=300 + i, flags=3) for i in range(P * rounds)]
    images = [s3.Image.from_array(v, units) for v in vols]
    for im in images:
        im.make_resident(0)
    dets = [s3.Detector(0) for _ in range(P)]
    kps = [s3.KeypointStore() for _ in range(P)]
    dss = [s3.DescriptorStore() for _ in range(P)]
    n = P * rounds
    s3.detect_describe_batch(dets, images, kps=[kps[v % P] for v in range(n)], dss=[dss[v % P] for v in range(n)])
    for p in range(P):
        keys, desc = _single(vols[(rounds - 1) * P + p], units)
        got = kps[p].records()
        assert len(got) == len(keys) and len(keys) > 0
        for f in keys.dtype.names:
            assert np.array_equal(got[f], keys[f]), (p, f)
        assert np.array_equal(dss[p].to_matrix().view(np.uint32), desc.view(np.uint32)), p
    for d in dets:
        d.close()
