"""ctypes wrapper of oracle/_ref/libsift3d_refstage.so -- the UNMODIFIED reference
compiled from /root/reference (oracle/Makefile) plus oracle/ref_stage_harness.c.

TEST INFRASTRUCTURE ONLY: importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs; never from the product.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REFSTAGE_PATH = os.path.join(_HERE, "_ref", "libsift3d_refstage.so")
REF_PATH = os.path.join(_HERE, "_ref", "libsift3D_ref.so")

CAND_DTYPE = np.dtype([("o", "<i4"), ("s", "<i4"), ("x", "<f8"), ("y", "<f8"), ("z", "<f8"),
                       ("sd", "<f8"), ("strength", "<f4"), ("pad", "<i4")], align=True)
KEY_DTYPE = np.dtype([("o", "<i4"), ("s", "<i4"), ("x", "<f8"), ("y", "<f8"), ("z", "<f8"),
                      ("sd", "<f8"), ("strength", "<f4"), ("R", "<f4", (9,)), ("pad", "<i4")],
                     align=True)

_lib = None


def available() -> bool:
    return os.path.exists(REFSTAGE_PATH)


def load():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(REFSTAGE_PATH)
    vp = C.c_void_p
    for fn in ("sift3d_make_detector", "sift3d_make_keypoint_store", "sift3d_make_descriptor_store",
               "sift3d_make_mat_rm", "sift3d_mat_rm_data"):
        getattr(L, fn).restype = vp
    L.sift3d_make_image.restype = vp
    L.sift3d_make_image.argtypes = [C.c_int] * 4
    L.sift3d_image_data.restype = C.POINTER(C.c_float)
    L.sift3d_image_data.argtypes = [vp]
    for fn in ("sift3d_free_detector", "sift3d_free_keypoint_store", "sift3d_free_descriptor_store",
               "sift3d_free_image", "sift3d_free_mat_rm", "sift3d_mat_rm_data"):
        getattr(L, fn).argtypes = [vp]
    for fn in ("sift3d_detector_set_peak_thresh", "sift3d_detector_set_corner_thresh",
               "sift3d_detector_set_sigma_n", "sift3d_detector_set_sigma0"):
        getattr(L, fn).argtypes = [vp, C.c_double]
    L.sift3d_detector_set_num_kp_levels.argtypes = [vp, C.c_uint]
    L.sift3d_detect_keypoints.argtypes = [vp, vp, vp]
    L.sift3d_extract_descriptors.argtypes = [vp, vp, vp]
    L.sift3d_keypoint_store_to_mat_rm.argtypes = [vp, vp]
    L.sift3d_descriptor_store_to_mat_rm.argtypes = [vp, vp]
    L.sift3d_keypoint_store_save.argtypes = [C.c_char_p, vp]
    L.sift3d_descriptor_store_save.argtypes = [C.c_char_p, vp]
    L.sift3d_keypoint_store_sort_by_strength.argtypes = [vp, C.c_int]
    L.sift3d_keypoint_store_sort_by_strength.restype = None
    L.sift3d_mat_rm_dimensions.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.sift3d_mat_rm_dimensions.restype = None
    L.sift3d_mat_rm_type.argtypes = [vp]
    # harness
    L.refstage_image_set_units.argtypes = [vp, C.c_double, C.c_double, C.c_double]
    L.refstage_image_set_units.restype = None
    L.refstage_detect.argtypes = [vp, vp, vp, vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_double)]
    L.refstage_num_octaves.argtypes = [vp]
    L.refstage_num_levels.argtypes = [vp, C.c_int]
    L.refstage_level.restype = C.POINTER(C.c_float)
    L.refstage_level.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int),
                                 C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.refstage_filter.argtypes = [vp, C.c_int, C.POINTER(C.c_float), C.c_int, C.POINTER(C.c_double)]
    L.refstage_mesh.argtypes = [vp, vp, vp]
    L.refstage_mesh.restype = None
    L.refstage_num_keypoints.argtypes = [vp]
    L.refstage_keypoints.argtypes = [vp, vp, C.c_int]
    L.refstage_keypoints.restype = None
    L.refstage_set_keypoints.argtypes = [vp, vp, C.c_int]
    L.refstage_blur.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                C.c_double, vp, C.c_int]
    L.refstage_eigen3.argtypes = [vp, vp, vp]
    L.refstage_sizeof.restype = C.c_size_t
    L.refstage_sizeof.argtypes = [C.c_int]
    _lib = L
    return L


def _mat_to_numpy(L, mat):
    cols, rows = C.c_int(), C.c_int()
    L.sift3d_mat_rm_dimensions(mat, C.byref(cols), C.byref(rows))
    dtype = {0: np.float64, 1: np.float32, 2: np.int32}[int(L.sift3d_mat_rm_type(mat))]
    n = rows.value * cols.value
    if n == 0:
        return np.zeros((rows.value, cols.value), dtype=dtype)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(L.sift3d_mat_rm_data(mat))
    return np.frombuffer(buf, dtype=dtype).reshape(rows.value, cols.value).copy()


class RefRun:
    """One reference detector + image + stores; detect()/extract() replay the public
    call sequence (cli/kpSift3D.c:96-146) with stage capture."""

    def __init__(self, vol: np.ndarray, units=(1.0, 1.0, 1.0), params=None):
        self.L = load()
        L = self.L
        vol = np.ascontiguousarray(vol, dtype=np.float32)
        nz, ny, nx = vol.shape
        self.shape = (nx, ny, nz)
        self.det = L.sift3d_make_detector()
        self.kp = L.sift3d_make_keypoint_store()
        self.desc = L.sift3d_make_descriptor_store()
        self.im = L.sift3d_make_image(nx, ny, nz, 1)
        assert self.det and self.kp and self.desc and self.im
        np.ctypeslib.as_array(L.sift3d_image_data(self.im), shape=(vol.size,))[:] = vol.ravel()
        L.refstage_image_set_units(self.im, *[float(u) for u in units])
        for k, v in (params or {}).items():
            rc = getattr(L, f"sift3d_detector_set_{k}")(self.det, v)
            assert rc == 0, f"reference rejected {k}={v}"
        self.candidates = None
        self.times = None

    def set_image(self, vol: np.ndarray, units=(1.0, 1.0, 1.0)):
        """Replace the image; the detector (and whatever it caches between images,
        sift.c:651-656) is kept, so detector reuse can be replayed."""
        L = self.L
        vol = np.ascontiguousarray(vol, dtype=np.float32)
        nz, ny, nx = vol.shape
        L.sift3d_free_image(self.im)
        self.im = L.sift3d_make_image(nx, ny, nz, 1)
        assert self.im
        self.shape = (nx, ny, nz)
        np.ctypeslib.as_array(L.sift3d_image_data(self.im), shape=(vol.size,))[:] = vol.ravel()
        L.refstage_image_set_units(self.im, *[float(u) for u in units])

    def detect(self, max_cand=1 << 22):
        L = self.L
        cand = np.zeros(max_cand, dtype=CAND_DTYPE)
        n = C.c_int(0)
        times = (C.c_double * 5)()
        rc = L.refstage_detect(self.det, self.im, self.kp, cand.ctypes.data_as(C.c_void_p),
                               max_cand, C.byref(n), times)
        if rc:
            raise RuntimeError("reference detect failed")
        assert n.value <= max_cand
        self.candidates = cand[: n.value].copy()
        self.times = dict(zip(("set_im", "gpyr", "dog", "extrema", "orient"), list(times)))
        return self.keypoints()

    def detect_public(self):
        """The public entry point, unstaged (what a user of the reference calls)."""
        if self.L.sift3d_detect_keypoints(self.det, self.im, self.kp):
            raise RuntimeError("reference detect failed")
        return self.keypoints()

    def keypoints(self) -> np.ndarray:
        n = self.L.refstage_num_keypoints(self.kp)
        out = np.zeros(n, dtype=KEY_DTYPE)
        if n:
            self.L.refstage_keypoints(self.kp, out.ctypes.data_as(C.c_void_p), n)
        return out

    def set_keypoints(self, keys: np.ndarray):
        keys = np.ascontiguousarray(keys, dtype=KEY_DTYPE)
        if self.L.refstage_set_keypoints(self.kp, keys.ctypes.data_as(C.c_void_p), len(keys)):
            raise RuntimeError("refstage_set_keypoints failed")

    def sort_by_strength(self, limit=0):
        self.L.sift3d_keypoint_store_sort_by_strength(self.kp, int(limit))

    def extract(self) -> np.ndarray:
        """N x 771 float32 descriptor matrix."""
        L = self.L
        if L.sift3d_extract_descriptors(self.det, self.kp, self.desc):
            raise RuntimeError("reference extract failed")
        mat = L.sift3d_make_mat_rm()
        try:
            if L.sift3d_descriptor_store_to_mat_rm(self.desc, mat):
                raise RuntimeError("reference to_mat_rm failed")
            return _mat_to_numpy(L, mat)
        finally:
            L.sift3d_free_mat_rm(mat)

    def keypoint_matrix(self) -> np.ndarray:
        L = self.L
        mat = L.sift3d_make_mat_rm()
        try:
            if L.sift3d_keypoint_store_to_mat_rm(self.kp, mat):
                raise RuntimeError("reference to_mat_rm failed")
            return _mat_to_numpy(L, mat)
        finally:
            L.sift3d_free_mat_rm(mat)

    def num_octaves(self):
        return self.L.refstage_num_octaves(self.det)

    def level(self, o, s, dog=False):
        dims = (C.c_int * 3)()
        units = (C.c_double * 3)()
        scale = C.c_double()
        p = self.L.refstage_level(self.det, int(dog), o, s, dims, units, C.byref(scale))
        if not p:
            raise IndexError((o, s))
        n = dims[0] * dims[1] * dims[2]
        arr = np.ctypeslib.as_array(p, shape=(n,)).reshape(dims[2], dims[1], dims[0]).copy()
        return arr, tuple(units), scale.value

    def filter(self, idx):
        taps = (C.c_float * 257)()
        sigma = C.c_double()
        w = self.L.refstage_filter(self.det, idx, taps, 257, C.byref(sigma))
        return np.array(taps[:w], dtype=np.float32), sigma.value

    def mesh(self):
        v = np.zeros((20, 3, 3), dtype=np.float32)
        idx = np.zeros((20, 3), dtype=np.int32)
        self.L.refstage_mesh(self.det, v.ctypes.data_as(C.c_void_p), idx.ctypes.data_as(C.c_void_p))
        return v, idx

    def close(self):
        L = self.L
        if self.det:
            L.sift3d_free_detector(self.det)
            L.sift3d_free_keypoint_store(self.kp)
            L.sift3d_free_descriptor_store(self.desc)
            L.sift3d_free_image(self.im)
            self.det = self.kp = self.desc = self.im = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def ref_blur(vol: np.ndarray, units, taps) -> np.ndarray:
    L = load()
    vol = np.ascontiguousarray(vol, dtype=np.float32)
    taps = np.ascontiguousarray(taps, dtype=np.float32)
    out = np.empty_like(vol)
    nz, ny, nx = vol.shape
    rc = L.refstage_blur(vol.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p), nx, ny, nz,
                         float(units[0]), float(units[1]), float(units[2]),
                         taps.ctypes.data_as(C.c_void_p), len(taps))
    if rc:
        raise RuntimeError("refstage_blur failed")
    return out


def ref_eigen3(A: np.ndarray):
    L = load()
    A = np.ascontiguousarray(A, dtype=np.float64)
    Q = np.zeros((3, 3))
    w = np.zeros(3)
    if L.refstage_eigen3(A.ctypes.data_as(C.c_void_p), Q.ctypes.data_as(C.c_void_p),
                         w.ctypes.data_as(C.c_void_p)):
        raise RuntimeError("refstage_eigen3 failed")
    return w, Q
