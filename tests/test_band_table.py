"""CPU test of the host logic behind the fast-mode blur for general tap spacings (blur_band.cuh):
the banded weight tables, applied in numpy, against the reference's own apply_Sep_FIR_filter
(oracle/_ref, imutil.c:680-861) -- every boundary rule, the carried coordinate of the interior
pass and the z-slab cut (org > 0) live in the table, so this pins them without a GPU."""
import ctypes as C

import numpy as np
import pytest

from oracle import refstage
from sift3d_b200 import api

needs_ref = pytest.mark.skipif(not refstage.available(), reason="oracle/_ref not built")


def _gauss_taps(sigma):
    hw = max(int(np.ceil(sigma * 3.0)), 1)
    x = (np.arange(2 * hw + 1, dtype=np.float64) - hw) / (sigma + np.finfo(np.float64).eps)
    k = np.exp(-0.5 * x * x).astype(np.float32)
    acc = np.float32(0)
    for v in k:
        acc = np.float32(acc + v)
    return (k / acc).astype(np.float32)


def band_table(n_line, org, unit, taps):
    L = api.libcuda()
    fn = L.sift3d_cuda_blur_band_table
    fn.argtypes = [C.c_int, C.c_int, C.c_float, C.POINTER(api.CudaFilter), C.POINTER(C.c_float), C.c_size_t,
                   C.POINTER(C.c_int), C.POINTER(C.c_int)]
    fn.restype = C.c_int
    f = api.CudaFilter()
    f.width = len(taps)
    for i, t in enumerate(taps):
        f.taps[i] = float(t)
    n = n_line - org
    cap = n * min(n, 4 * len(taps) + 8)
    w = (C.c_float * cap)()
    base = (C.c_int * n)()
    band = C.c_int(0)
    rc = fn(n_line, org, np.float32(1.0 / unit), C.byref(f), w, cap, base, C.byref(band))
    assert rc == 0, rc
    B = band.value
    return np.frombuffer(w, dtype=np.float32, count=B * n).reshape(B, n).copy(), np.array(base), B


def apply_band(vol, axis, w, base, B):
    """out[p] = sum_j w[j, p] * in[base[p] + j] along `axis` (float64 accumulation)."""
    v = np.moveaxis(vol.astype(np.float64), axis, 0)
    out = np.zeros_like(v)
    n = v.shape[0]
    for p in range(n):
        out[p] = np.tensordot(w[:, p].astype(np.float64), v[base[p]:base[p] + B], axes=(0, 0))
    return np.moveaxis(out, 0, axis)


@needs_ref
@pytest.mark.parametrize("units", [(1, 1, 1), (2, 2, 2), (4, 4, 4), (0.75, 0.75, 1.5), (1.0, 2.0, 0.5),
                                   (0.9, 1.3, 0.7), (3.0, 1.0, 8.0)])
@pytest.mark.parametrize("sigma", [0.5387, 1.2263, 2.4525])
def test_band_tables_reproduce_the_reference_blur(units, sigma):
    rng = np.random.default_rng(int(1000 * sigma) + int(8 * units[0]))
    nz, ny, nx = 29, 24, 41
    vol = rng.standard_normal((nz, ny, nx)).astype(np.float32)
    taps = _gauss_taps(sigma)
    want = refstage.ref_blur(vol, units, taps)
    cur = vol
    for axis, n, unit in ((2, nx, units[0]), (1, ny, units[1]), (0, nz, units[2])):
        w, base, B = band_table(n, 0, unit, taps)
        assert base.min() >= 0 and (base + B).max() <= n
        hw = len(taps) // 2
        assert B <= min(n, 2 * int(np.ceil(hw / unit)) + 3)
        # the passes of the reference store float32 between axes
        cur = apply_band(cur, axis, w, base, B).astype(np.float32)
    err = np.abs(cur.astype(np.float64) - want).max() / np.abs(want).max()
    assert err <= 5e-7, err


@needs_ref
@pytest.mark.parametrize("org,top", [(0, 40), (9, 40), (9, 31), (17, 57)])
def test_band_table_of_a_slab_cut(org, top):
    """z pass of a slab holding planes [org, top) of a 57-plane volume: the rows of the table whose
    reach stays inside the slab equal the rows of the whole line's table (shifted by org)."""
    n_vol, unit, taps = 57, 2.0, _gauss_taps(1.5450)
    hw = len(taps) // 2
    reach = int(np.ceil(hw / unit)) + 2
    wv, bv, Bv = band_table(n_vol, 0, unit, taps)
    # the product describes a slab by the line it is cut from: n_line = planes up to its top
    ws, bs, Bs = band_table(top, org, unit, taps)
    lo = org + (reach if org > 0 else 0)
    hi = top - (reach if top < n_vol else 0)
    for p in range(lo, hi):
        full = np.zeros(n_vol)
        full[bv[p]:bv[p] + Bv] = wv[:, p]
        cut = np.zeros(n_vol)
        b = bs[p - org] + org
        cut[b:b + Bs] = ws[:, p - org]
        assert np.array_equal(full, cut), p


def test_band_table_rejects_bad_arguments():
    L = api.libcuda()
    f = api.CudaFilter()
    f.width = 4  # even
    w = (C.c_float * 16)()
    b = (C.c_int * 4)()
    band = C.c_int(0)
    L.sift3d_cuda_blur_band_table.restype = C.c_int
    L.sift3d_cuda_blur_band_table.argtypes = [C.c_int, C.c_int, C.c_float, C.POINTER(api.CudaFilter),
                                              C.POINTER(C.c_float), C.c_size_t, C.POINTER(C.c_int),
                                              C.POINTER(C.c_int)]
    assert L.sift3d_cuda_blur_band_table(4, 0, 1.0, C.byref(f), w, 16, b, C.byref(band)) == -1
    f.width = 3
    assert L.sift3d_cuda_blur_band_table(1, 0, 1.0, C.byref(f), w, 16, b, C.byref(band)) == -1
    assert L.sift3d_cuda_blur_band_table(8, 0, 0.0, C.byref(f), w, 16, b, C.byref(band)) == -1
    # capacity too small for the band
    f.taps[0], f.taps[1], f.taps[2] = 0.25, 0.5, 0.25
    assert L.sift3d_cuda_blur_band_table(8, 0, 1.0, C.byref(f), w, 2, b, C.byref(band)) == 1
