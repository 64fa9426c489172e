"""Deterministic synthetic workloads (BASELINE.json `configs`), generated
by libsift3d_synth.so (sift3d_b200/synth.c)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from .api import Sift3DError, _load

DEFAULT_SEED = 88172645463325252

# name -> (nx, ny, nz), units, blobs, flags (bit0 bias field, bit1 noise)
WORKLOADS = {
    "C1_blobs_128": ((128, 128, 128), (1.0, 1.0, 1.0), 128 ** 3 // 4096, 0),
    "C2_mri_256": ((256, 256, 256), (1.0, 1.0, 1.0), 256 ** 3 // 4096, 3),
    "C3_ct_512x512x300": ((512, 512, 300), (0.75, 0.75, 1.5), 512 * 512 * 300 // 4096, 0),
    "C5_blobs_1024": ((1024, 1024, 1024), (1.0, 1.0, 1.0), 1024 ** 3 // 4096, 0),
}


def synth_volume(nx, ny, nz, num_blobs=None, seed=DEFAULT_SEED, flags=0, out=None) -> np.ndarray:
    """(nz, ny, nx) float32 blob volume; `out` may be a preallocated array (e.g. the
    page-locked buffer of an Image)."""
    L = _load("synth")
    L.sift3d_synth_volume.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                      C.c_int]
    if num_blobs is None:
        num_blobs = max(1, nx * ny * nz // 4096)
    if out is None:
        out = np.empty((nz, ny, nx), dtype=np.float32)
    assert out.dtype == np.float32 and out.size == nx * ny * nz and out.flags.c_contiguous
    if L.sift3d_synth_volume(out.ctypes.data_as(C.c_void_p), nx, ny, nz, int(num_blobs),
                             C.c_uint64(seed), int(flags)):
        raise Sift3DError("sift3d_synth_volume failed")
    return out


def workload(name, seed=DEFAULT_SEED, out=None):
    """Returns (volume, units) of a named configuration."""
    (nx, ny, nz), units, blobs, flags = WORKLOADS[name]
    return synth_volume(nx, ny, nz, blobs, seed, flags, out), units
