"""ctypes wrapper of oracle/libsift3d_oracle.so -- our CPU restatement of the reference's
detect + describe path (oracle/sift3d_oracle.c). TEST INFRASTRUCTURE ONLY."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .refstage import CAND_DTYPE, KEY_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
PORT_PATH = os.path.join(_HERE, "libsift3d_oracle.so")
_lib = None


def available() -> bool:
    return os.path.exists(PORT_PATH)


def load():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(PORT_PATH)
    vp = C.c_void_p
    L.orc_make.restype = vp
    L.orc_free.argtypes = [vp]
    L.orc_set_params.argtypes = [vp, C.c_double, C.c_double, C.c_int, C.c_double, C.c_double]
    L.orc_detect.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double]
    L.orc_num_octaves.argtypes = [vp]
    L.orc_num_levels.argtypes = [vp]
    L.orc_level.restype = C.POINTER(C.c_float)
    L.orc_level.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double),
                            C.POINTER(C.c_double)]
    L.orc_filter.argtypes = [vp, C.c_int, C.POINTER(C.c_float), C.c_int]
    L.orc_num_candidates.argtypes = [vp]
    L.orc_candidates.argtypes = [vp, vp, C.c_int]
    L.orc_candidates.restype = None
    L.orc_num_keypoints.argtypes = [vp]
    L.orc_keypoints.argtypes = [vp, vp, C.c_int]
    L.orc_keypoints.restype = None
    L.orc_sort_by_strength.argtypes = [vp, C.c_int]
    L.orc_sort_by_strength.restype = None
    L.orc_describe.argtypes = [vp, vp, C.c_int, vp]
    L.orc_blur.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double, C.c_double,
                           vp, C.c_int]
    L.orc_mesh.argtypes = [vp, vp]
    L.orc_mesh.restype = None
    L.orc_eigen3.argtypes = [vp, vp, vp]
    L.orc_eigen3.restype = None
    _lib = L
    return L


class PortRun:
    """Same surface as refstage.RefRun, backed by the restatement."""

    DEFAULTS = dict(peak_thresh=0.1, corner_thresh=0.4, num_kp_levels=3, sigma_n=1.15, sigma0=1.6)

    def __init__(self, vol: np.ndarray, units=(1.0, 1.0, 1.0), params=None):
        self.L = load()
        self.vol = np.ascontiguousarray(vol, dtype=np.float32)
        self.units = tuple(float(u) for u in units)
        self.det = self.L.orc_make()
        p = dict(self.DEFAULTS)
        p.update(params or {})
        if self.L.orc_set_params(self.det, p["peak_thresh"], p["corner_thresh"],
                                 int(p["num_kp_levels"]), p["sigma_n"], p["sigma0"]):
            raise RuntimeError("oracle rejected the parameters")
        self.candidates = None

    def detect(self):
        nz, ny, nx = self.vol.shape
        if self.L.orc_detect(self.det, self.vol.ctypes.data_as(C.c_void_p), nx, ny, nz, *self.units):
            raise RuntimeError("oracle detect failed")
        n = self.L.orc_num_candidates(self.det)
        self.candidates = np.zeros(n, dtype=CAND_DTYPE)
        if n:
            self.L.orc_candidates(self.det, self.candidates.ctypes.data_as(C.c_void_p), n)
        return self.keypoints()

    def keypoints(self):
        n = self.L.orc_num_keypoints(self.det)
        out = np.zeros(n, dtype=KEY_DTYPE)
        if n:
            self.L.orc_keypoints(self.det, out.ctypes.data_as(C.c_void_p), n)
        return out

    def sort_by_strength(self, limit=0):
        self.L.orc_sort_by_strength(self.det, int(limit))

    def describe(self, keys: np.ndarray) -> np.ndarray:
        keys = np.ascontiguousarray(keys, dtype=KEY_DTYPE)
        out = np.zeros((len(keys), 768), dtype=np.float32)
        if self.L.orc_describe(self.det, keys.ctypes.data_as(C.c_void_p), len(keys),
                               out.ctypes.data_as(C.c_void_p)):
            raise RuntimeError("oracle describe failed")
        return out

    def extract(self) -> np.ndarray:
        """N x 771 like sift3d_descriptor_store_to_mat_rm (sift.c:1683-1726)."""
        keys = self.keypoints()
        d = self.describe(keys)
        f = np.ldexp(1.0, keys["o"])
        xyz = np.stack([keys["x"] * f, keys["y"] * f, keys["z"] * f], axis=1).astype(np.float32)
        return np.concatenate([xyz, d], axis=1)

    def keypoint_matrix(self) -> np.ndarray:
        keys = self.keypoints()
        f = np.ldexp(1.0, keys["o"])
        return np.stack([keys["x"] * f, keys["y"] * f, keys["z"] * f], axis=1)

    def num_octaves(self):
        return self.L.orc_num_octaves(self.det)

    def num_levels(self):
        return self.L.orc_num_levels(self.det)

    def level(self, o, s):
        dims = (C.c_int * 3)()
        units = (C.c_double * 3)()
        scale = C.c_double()
        p = self.L.orc_level(self.det, o, s, dims, units, C.byref(scale))
        if not p:
            raise IndexError((o, s))
        n = dims[0] * dims[1] * dims[2]
        arr = np.ctypeslib.as_array(p, shape=(n,)).reshape(dims[2], dims[1], dims[0]).copy()
        return arr, tuple(units), scale.value

    def filter(self, idx):
        taps = (C.c_float * 513)()
        w = self.L.orc_filter(self.det, idx, taps, 513)
        return np.array(taps[:w], dtype=np.float32)

    def close(self):
        if self.det:
            self.L.orc_free(self.det)
            self.det = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def port_blur(vol, units, taps):
    L = load()
    vol = np.ascontiguousarray(vol, dtype=np.float32)
    taps = np.ascontiguousarray(taps, dtype=np.float32)
    out = np.empty_like(vol)
    nz, ny, nx = vol.shape
    if L.orc_blur(vol.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), nx, ny, nz,
                  float(units[0]), float(units[1]), float(units[2]),
                  taps.ctypes.data_as(C.c_void_p), len(taps)):
        raise RuntimeError("orc_blur failed")
    return out


def port_mesh():
    L = load()
    v = np.zeros((20, 3, 3), dtype=np.float32)
    idx = np.zeros((20, 3), dtype=np.int32)
    L.orc_mesh(v.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p))
    return v, idx


def port_eigen3(A):
    L = load()
    A = np.ascontiguousarray(A, dtype=np.float64)
    Q = np.zeros((3, 3))
    w = np.zeros(3)
    L.orc_eigen3(A.ctypes.data_as(C.c_void_p), Q.ctypes.data_as(C.c_void_p),
                 w.ctypes.data_as(C.c_void_p))
    return w, Q
