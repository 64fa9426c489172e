"""sift3d_cuda_blur_static: the blur entry point that allocates nothing (its scratch is a
file-scope __device__ array of libsift3d_cuda.so; include/sift3d_cuda.h). No counterpart in the
reference. Checked through the C ABI against
the existing entry point sift3d_cuda_blur on two calls in a row (the second one smaller, so
stale scratch contents of the first would show). Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

import parity

pytestmark = pytest.mark.gpu

MAX_VOX = 1 << 20  # SIFT3D_CUDA_STATIC_BLUR_MAX_VOXELS


def _gauss_taps(sigma):
    hw = max(int(np.ceil(sigma * 3.0)), 1)
    x = (np.arange(2 * hw + 1, dtype=np.float64) - hw) / (sigma + np.finfo(np.float64).eps)
    k = np.exp(-0.5 * x * x).astype(np.float32)
    acc = np.float32(0)
    for v in k:
        acc = np.float32(acc + v)
    return (k / acc).astype(np.float32)


def _blur_static(L, device, vol, units, taps, exact):
    """Input in device memory (sift3d_cuda_device_upload), output in page-locked host memory
    written by the kernels through its unified address; returns (rc, array)."""
    from sift3d_b200.api import CudaFilter
    nz, ny, nx = vol.shape
    f = CudaFilter()
    f.width = len(taps)
    for i, t in enumerate(taps):
        f.taps[i] = float(t)
    u = (C.c_double * 3)(*[float(x) for x in units])
    d_src = L.sift3d_cuda_device_upload(device, vol.ctypes.data_as(C.c_void_p), vol.size)
    assert d_src
    h_dst = L.sift3d_cuda_host_alloc(vol.nbytes)
    assert h_dst
    try:
        out = np.ctypeslib.as_array(C.cast(h_dst, C.POINTER(C.c_float)), shape=(vol.size,))
        out[:] = np.nan
        rc = L.sift3d_cuda_blur_static(device, d_src, h_dst, nx, ny, nz, u, C.byref(f), int(exact),
                                       None, 1)
        return rc, out.reshape(vol.shape).copy()
    finally:
        L.sift3d_cuda_device_free(device, d_src)
        L.sift3d_cuda_host_free(h_dst)


@pytest.mark.parametrize("exact", [True, False], ids=["exact", "fast"])
def test_blur_static_matches_blur_two_calls_in_a_row(exact):
    import sift3d_b200 as s3
    L = s3.libcuda()
    det = s3.Detector()
    det.set_exact(exact)
    dev = det.device()
    rng = np.random.default_rng(20261105)
    # (dims (nz, ny, nx), units, sigma): generic 3-pass path (uses both scratch volumes), then a
    # smaller fused-eligible volume, then generic again on a different shape
    calls = [((40, 64, 96), (0.75, 0.75, 1.5), 1.2263),
             ((36, 40, 64), (1.0, 1.0, 1.0), 0.9733),
             ((23, 20, 37), (2.0, 1.0, 0.5), 1.5450)]
    try:
        for shape, units, sigma in calls:
            vol = rng.standard_normal(shape).astype(np.float32)
            taps = _gauss_taps(sigma)
            want = det.blur(vol, units, taps)
            rc, got = _blur_static(L, dev, vol, units, taps, exact)
            assert rc == 0
            # exact mode and the fused kernel: the same kernels, 0 differing bits (the tolerance of
            # the exact-mode tests). Fast mode with a general spacing: the existing entry point
            # runs the banded kernels (their tables are allocated at first use), this one the
            # literal kernels -- the fast-mode tolerance of tests/test_gpu_parity.py
            rel, nbits = parity.level_error(got, want)
            if exact or units == (1.0, 1.0, 1.0):
                assert nbits == 0, f"{shape} {units}: {nbits} voxels differ (max rel err {rel:.3g})"
            else:
                assert rel <= 2e-6, f"{shape} {units}: max rel err {rel:.3g}"
    finally:
        det.close()


def test_blur_static_rejects_oversize_and_bad_arguments():
    import sift3d_b200 as s3
    L = s3.libcuda()
    from sift3d_b200.api import CudaFilter
    f = CudaFilter()
    f.width = 3
    f.taps[0], f.taps[1], f.taps[2] = 0.25, 0.5, 0.25
    u = (C.c_double * 3)(1.0, 1.0, 1.0)
    buf = L.sift3d_cuda_host_alloc(64)
    try:
        # 128 x 128 x 65 voxels > 2^20: rejected before anything is touched
        assert 128 * 128 * 65 > MAX_VOX
        assert L.sift3d_cuda_blur_static(0, buf, C.c_void_p(buf + 32), 128, 128, 65, u, C.byref(f), 0, None, 1) == -1
        # in place, even width, NULL
        assert L.sift3d_cuda_blur_static(0, buf, buf, 8, 8, 8, u, C.byref(f), 0, None, 1) == -1
        f.width = 4
        assert L.sift3d_cuda_blur_static(0, buf, C.c_void_p(buf + 32), 2, 2, 2, u, C.byref(f), 0, None, 1) == -1
        f.width = 3
        assert L.sift3d_cuda_blur_static(0, None, buf, 8, 8, 8, u, C.byref(f), 0, None, 1) == -1
    finally:
        L.sift3d_cuda_host_free(buf)
