"""ctypes bindings of libsift3D.so (include/sift3d.h) and libsift3d_cuda.so
(include/sift3d_cuda.h). Mirrors the reference's user-facing call sequence
(doc/Manual.md:34-77, cli/kpSift3D.c:96-146): make image -> fill data -> make
detector/stores -> detect -> (sort) -> extract -> to_mat_rm / save.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIBDIR = os.path.join(_HERE, "lib")

SIFT3D_SUCCESS = 0
SIFT3D_FAILURE = -1
DESC_NUMEL = 768
MAX_FILTERS = 16
MAX_OCTAVES = 24
MAX_TAPS = 129


class Sift3DError(RuntimeError):
    pass


def lib_paths():
    return {
        "host": os.path.join(_LIBDIR, "libsift3D.so"),
        "cuda": os.path.join(_LIBDIR, "libsift3d_cuda.so"),
        "synth": os.path.join(_LIBDIR, "libsift3d_synth.so"),
    }


class CudaFilter(C.Structure):
    _fields_ = [("width", C.c_int), ("taps", C.c_float * MAX_TAPS)]


class CudaPlanDesc(C.Structure):
    _fields_ = [
        ("nx", C.c_int), ("ny", C.c_int), ("nz", C.c_int), ("units", C.c_double * 3),
        ("num_octaves", C.c_int), ("num_kp_levels", C.c_int),
        ("scales", (C.c_double * (MAX_FILTERS + 1)) * MAX_OCTAVES),
        ("first", CudaFilter), ("octave", CudaFilter * MAX_FILTERS),
    ]


class CudaSlabDesc(C.Structure):
    _fields_ = [("num_sharded_octaves", C.c_int), ("own0", C.c_int * MAX_OCTAVES),
                ("own1", C.c_int * MAX_OCTAVES), ("org", C.c_int * MAX_OCTAVES),
                ("vol_nz", C.c_int * MAX_OCTAVES)]


REDUCE_MAX_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_void_p, C.c_int)
EXCHANGE_FN = C.CFUNCTYPE(C.c_int, C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_int,
                          C.c_int, C.c_int, C.c_int)


class CudaSlabCallbacks(C.Structure):
    _fields_ = [("user", C.c_void_p), ("reduce_max", REDUCE_MAX_FN), ("exchange", EXCHANGE_FN),
                ("in_stream", C.c_int), ("input_reach", C.c_int)]


class CudaKeypoint(C.Structure):
    _fields_ = [
        ("o", C.c_int32), ("s", C.c_int32),
        ("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("sd", C.c_double),
        ("strength", C.c_float), ("R", C.c_float * 9), ("valid", C.c_int32),
    ]


class CudaTimings(C.Structure):
    _fields_ = [(n, C.c_float) for n in (
        "h2d", "scale", "pyramid", "dogmax", "extrema", "order", "orient", "d2h", "total_detect",
        "desc_h2d", "descriptors", "desc_d2h", "total_describe")]

    def as_dict(self):
        return {n: float(getattr(self, n)) for n, _ in self._fields_}


# numpy view of sift3d_cuda_keypoint
CUDA_KEYPOINT_DTYPE = np.dtype([
    ("o", "<i4"), ("s", "<i4"), ("x", "<f8"), ("y", "<f8"), ("z", "<f8"), ("sd", "<f8"),
    ("strength", "<f4"), ("R", "<f4", (9,)), ("valid", "<i4")], align=True)
assert CUDA_KEYPOINT_DTYPE.itemsize == C.sizeof(CudaKeypoint)

_libs = {}


def _load(name):
    if name in _libs:
        return _libs[name]
    path = lib_paths()[name]
    if not os.path.exists(path):
        raise Sift3DError(
            f"{path} is missing: build it with `make product` (python __graft_entry__.py build). "
            "There is no Python/CPU fallback for the CUDA path.")
    mode = C.RTLD_GLOBAL if name == "cuda" else C.DEFAULT_MODE
    handle = C.CDLL(path, mode=mode)
    _libs[name] = handle
    return handle


def libcuda():
    """libsift3d_cuda.so with argument types declared."""
    L = _load("cuda")
    if getattr(L, "_typed", False):
        return L
    vp = C.c_void_p
    L.sift3d_cuda_device_count.restype = C.c_int
    L.sift3d_cuda_ctx_create.restype = vp
    L.sift3d_cuda_ctx_create.argtypes = [C.c_int]
    L.sift3d_cuda_ctx_destroy.argtypes = [vp]
    L.sift3d_cuda_host_alloc.restype = vp
    L.sift3d_cuda_host_alloc.argtypes = [C.c_size_t]
    L.sift3d_cuda_host_free.argtypes = [vp]
    L.sift3d_cuda_detect.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_double), C.c_double,
                                     C.c_double, C.POINTER(C.c_int)]
    L.sift3d_cuda_plan.argtypes = [vp, C.POINTER(CudaPlanDesc)]
    L.sift3d_cuda_detect_slab.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_double), C.c_double,
                                          C.c_double, C.POINTER(CudaSlabDesc),
                                          C.POINTER(CudaSlabCallbacks), C.POINTER(C.c_int)]
    L.sift3d_cuda_detect_from_seed.argtypes = [vp, vp, C.c_int, C.POINTER(C.c_double),
                                               C.c_double, C.c_double, C.POINTER(C.c_int)]
    L.sift3d_cuda_set_base_octave.argtypes = [vp, C.c_int]
    L.sift3d_cuda_level_ptr.argtypes = [vp, C.c_int, C.c_int]
    L.sift3d_cuda_level_ptr.restype = vp
    L.sift3d_cuda_fetch_candidates.argtypes = [vp, vp, C.c_int]
    L.sift3d_cuda_describe.argtypes = [vp, vp, C.c_int, vp, C.c_int]
    L.sift3d_cuda_level_info.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_int),
                                         C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.sift3d_cuda_fetch_level.argtypes = [vp, C.c_int, C.c_int, vp]
    L.sift3d_cuda_dogmax.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float)]
    L.sift3d_cuda_blur.argtypes = [vp, vp, vp, C.c_int, C.c_int, C.c_int, C.POINTER(C.c_double),
                                   C.POINTER(CudaFilter), C.c_int]
    L.sift3d_cuda_get_timings.argtypes = [vp, C.POINTER(CudaTimings)]
    L.sift3d_cuda_blur_static.argtypes = [C.c_int, vp, vp, C.c_int, C.c_int, C.c_int,
                                          C.POINTER(C.c_double), C.POINTER(CudaFilter), C.c_int, vp, C.c_int]
    L.sift3d_cuda_device_upload.restype = vp
    L.sift3d_cuda_device_upload.argtypes = [C.c_int, vp, C.c_size_t]
    L.sift3d_cuda_device_free.argtypes = [C.c_int, vp]
    L.sift3d_cuda_launch_count.restype = C.c_uint64
    L.sift3d_cuda_launch_count.argtypes = [vp]
    L.sift3d_cuda_synchronize.argtypes = [vp]
    L.sift3d_cuda_ctx_device.argtypes = [vp]
    L.sift3d_cuda_mark.argtypes = [vp, C.c_int]
    L.sift3d_cuda_elapsed_ms.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float)]
    L.sift3d_cuda_set_profiling.argtypes = [vp, C.c_int]
    L.sift3d_cuda_set_exact.argtypes = [vp, C.c_int]
    L.sift3d_cuda_get_exact.argtypes = [vp]
    L.sift3d_cuda_blur_ms.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_float),
                                      C.POINTER(C.c_int)]
    L._typed = True
    return L


def lib():
    """libsift3D.so (the drop-in public API) with argument types declared."""
    libcuda()  # dependency, loaded RTLD_GLOBAL first so $ORIGIN lookup is not needed
    L = _load("host")
    if getattr(L, "_typed", False):
        return L
    vp = C.c_void_p
    for fn in ("sift3d_make_detector", "sift3d_make_keypoint_store", "sift3d_make_descriptor_store",
               "sift3d_make_mat_rm", "sift3d_mat_rm_data", "sift3d_detector_cuda_ctx"):
        getattr(L, fn).restype = vp
    L.sift3d_make_detector_on_device.restype = vp
    L.sift3d_make_detector_on_device.argtypes = [C.c_int]
    L.sift3d_make_image.restype = vp
    L.sift3d_make_image.argtypes = [C.c_int] * 4
    L.sift3d_read_image.restype = vp
    L.sift3d_read_image.argtypes = [C.c_char_p]
    L.sift3d_image_data.restype = C.POINTER(C.c_float)
    L.sift3d_image_data.argtypes = [vp]
    for fn in ("sift3d_free_detector", "sift3d_free_keypoint_store", "sift3d_free_descriptor_store",
               "sift3d_free_image", "sift3d_free_mat_rm", "sift3d_mat_rm_data",
               "sift3d_detector_cuda_ctx"):
        getattr(L, fn).argtypes = [vp]
    for fn in ("sift3d_detector_set_peak_thresh", "sift3d_detector_set_corner_thresh",
               "sift3d_detector_set_sigma_n", "sift3d_detector_set_sigma0"):
        getattr(L, fn).argtypes = [vp, C.c_double]
    L.sift3d_detector_set_num_kp_levels.argtypes = [vp, C.c_uint]
    L.sift3d_detect_keypoints.argtypes = [vp, vp, vp]
    L.sift3d_extract_descriptors.argtypes = [vp, vp, vp]
    L.SIFT3D_detect_keypoints.argtypes = [vp, vp, vp]
    L.SIFT3D_extract_descriptors.argtypes = [vp, vp, vp]
    L.sift3d_detect_describe_batch.argtypes = [C.POINTER(vp), C.c_int, C.POINTER(vp), C.POINTER(vp),
                                               C.POINTER(vp), C.c_int]
    L.sift3d_detect_describe_sharded.argtypes = [vp, vp, C.POINTER(C.c_int), C.c_int, C.c_int, vp, vp]
    L.sift3d_sharded_timings.argtypes = [vp, C.POINTER(C.c_float)]
    L.sift3d_sharded_layout.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_double] + [C.POINTER(C.c_int)] * 4
    L.sift3d_keypoint_store_to_mat_rm.argtypes = [vp, vp]
    L.sift3d_descriptor_store_to_mat_rm.argtypes = [vp, vp]
    L.sift3d_keypoint_store_save.argtypes = [C.c_char_p, vp]
    L.sift3d_descriptor_store_save.argtypes = [C.c_char_p, vp]
    L.sift3d_keypoint_store_sort_by_strength.argtypes = [vp, C.c_int]
    L.sift3d_keypoint_store_sort_by_strength.restype = None
    L.sift3d_mat_rm_dimensions.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.sift3d_mat_rm_dimensions.restype = None
    L.sift3d_mat_rm_type.argtypes = [vp]
    L.sift3d_image_set_units.argtypes = [vp, C.c_double, C.c_double, C.c_double]
    L.sift3d_detector_fill_plan.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_double, C.c_double,
                                            C.c_double, C.c_int, C.c_int, C.POINTER(CudaPlanDesc)]
    L.sift3d_image_get_units.argtypes = [vp] + [C.POINTER(C.c_double)] * 3
    L.sift3d_image_get_units.restype = None
    L.sift3d_image_dimensions.argtypes = [vp] + [C.POINTER(C.c_int)] * 4
    L.sift3d_image_dimensions.restype = None
    L.sift3d_image_make_resident.argtypes = [vp, C.c_int]
    L.sift3d_image_prefetch.argtypes = [vp, C.c_int]
    L.sift3d_image_prefetch_wait.argtypes = [vp]
    L.sift3d_keypoint_store_count.argtypes = [vp]
    L.sift3d_descriptor_store_count.argtypes = [vp]
    L.sift3d_keypoint_store_get.argtypes = [vp, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int),
                                            C.POINTER(C.c_double), C.POINTER(C.c_double),
                                            C.POINTER(C.c_float), C.POINTER(C.c_float)]
    L._typed = True
    return L


def cuda_device_count() -> int:
    return int(libcuda().sift3d_cuda_device_count())


def mat_rm_to_numpy(L, mat) -> np.ndarray:
    """Copy a sift3d_mat_rm into a numpy array (rows x cols)."""
    cols, rows = C.c_int(), C.c_int()
    L.sift3d_mat_rm_dimensions(mat, C.byref(cols), C.byref(rows))
    t = L.sift3d_mat_rm_type(mat)
    dtype = {0: np.float64, 1: np.float32, 2: np.int32}[int(t)]
    n = rows.value * cols.value
    if n == 0:
        return np.zeros((rows.value, cols.value), dtype=dtype)
    ptr = L.sift3d_mat_rm_data(mat)
    buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
    return np.frombuffer(buf, dtype=dtype).reshape(rows.value, cols.value).copy()


class Image:
    """sift3d_image: nx*ny*nz float32, x fastest. ``array`` is a numpy view of the
    library-owned (page-locked) buffer with shape (nz, ny, nx)."""

    def __init__(self, nx, ny, nz, nc=1, units=(1.0, 1.0, 1.0)):
        self._L = lib()
        self.handle = self._L.sift3d_make_image(int(nx), int(ny), int(nz), int(nc))
        if not self.handle:
            raise Sift3DError("sift3d_make_image failed")
        self.nx, self.ny, self.nz, self.nc = int(nx), int(ny), int(nz), int(nc)
        ptr = self._L.sift3d_image_data(self.handle)
        n = self.nx * self.ny * self.nz * self.nc
        shape = (self.nz, self.ny, self.nx) if nc == 1 else (self.nz, self.ny, self.nx, self.nc)
        self.array = np.ctypeslib.as_array(ptr, shape=(n,)).reshape(shape)
        if tuple(units) != (1.0, 1.0, 1.0):
            self.set_units(*units)

    @classmethod
    def from_array(cls, vol: np.ndarray, units=(1.0, 1.0, 1.0)) -> "Image":
        vol = np.asarray(vol, dtype=np.float32)
        assert vol.ndim == 3, "expected a (nz, ny, nx) volume"
        im = cls(vol.shape[2], vol.shape[1], vol.shape[0], 1, units)
        im.array[...] = vol
        return im

    def make_resident(self, device=0):
        if self._L.sift3d_image_make_resident(self.handle, int(device)):
            raise Sift3DError("sift3d_image_make_resident failed")

    def prefetch(self, device=0):
        """Start the asynchronous upload that the next detect on `device` will consume."""
        if self._L.sift3d_image_prefetch(self.handle, int(device)):
            raise Sift3DError("sift3d_image_prefetch failed")

    def prefetch_wait(self):
        if self._L.sift3d_image_prefetch_wait(self.handle):
            raise Sift3DError("sift3d_image_prefetch_wait failed")

    def set_units(self, ux, uy, uz):
        if self._L.sift3d_image_set_units(self.handle, float(ux), float(uy), float(uz)):
            raise Sift3DError("sift3d_image_set_units failed")

    def close(self):
        if getattr(self, "handle", None):
            self.array = None
            self._L.sift3d_free_image(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class KeypointStore:
    def __init__(self):
        self._L = lib()
        self.handle = self._L.sift3d_make_keypoint_store()
        if not self.handle:
            raise Sift3DError("sift3d_make_keypoint_store failed")

    def __len__(self):
        return int(self._L.sift3d_keypoint_store_count(self.handle))

    def sort_by_strength(self, limit=0):
        self._L.sift3d_keypoint_store_sort_by_strength(self.handle, int(limit))

    def to_matrix(self) -> np.ndarray:
        mat = self._L.sift3d_make_mat_rm()
        try:
            if self._L.sift3d_keypoint_store_to_mat_rm(self.handle, mat):
                raise Sift3DError("sift3d_keypoint_store_to_mat_rm failed")
            return mat_rm_to_numpy(self._L, mat)
        finally:
            self._L.sift3d_free_mat_rm(mat)

    def save(self, path) -> int:
        return int(self._L.sift3d_keypoint_store_save(os.fsencode(path), self.handle))

    def records(self) -> np.ndarray:
        """All keypoints as a structured array (o, s, x, y, z, sd, strength, R[9])."""
        n = len(self)
        out = np.zeros(n, dtype=[("o", "i4"), ("s", "i4"), ("x", "f8"), ("y", "f8"), ("z", "f8"),
                                 ("sd", "f8"), ("strength", "f4"), ("R", "f4", (9,))])
        o, s = C.c_int(), C.c_int()
        xyz = (C.c_double * 3)()
        sd = C.c_double()
        st = C.c_float()
        R = (C.c_float * 9)()
        for i in range(n):
            self._L.sift3d_keypoint_store_get(self.handle, i, C.byref(o), C.byref(s), xyz,
                                              C.byref(sd), C.byref(st), R)
            out[i] = (o.value, s.value, xyz[0], xyz[1], xyz[2], sd.value, st.value, tuple(R))
        return out

    def close(self):
        if getattr(self, "handle", None):
            self._L.sift3d_free_keypoint_store(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class DescriptorStore:
    def __init__(self):
        self._L = lib()
        self.handle = self._L.sift3d_make_descriptor_store()
        if not self.handle:
            raise Sift3DError("sift3d_make_descriptor_store failed")

    def __len__(self):
        return int(self._L.sift3d_descriptor_store_count(self.handle))

    def to_matrix(self) -> np.ndarray:
        """N x 771 float32: x, y, z, 768 descriptor elements."""
        mat = self._L.sift3d_make_mat_rm()
        try:
            if self._L.sift3d_descriptor_store_to_mat_rm(self.handle, mat):
                raise Sift3DError("sift3d_descriptor_store_to_mat_rm failed")
            return mat_rm_to_numpy(self._L, mat)
        finally:
            self._L.sift3d_free_mat_rm(mat)

    def save(self, path) -> int:
        return int(self._L.sift3d_descriptor_store_save(os.fsencode(path), self.handle))

    def close(self):
        if getattr(self, "handle", None):
            self._L.sift3d_free_descriptor_store(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Detector:
    """sift3d_detector. Raises Sift3DError when no CUDA device is usable."""

    def __init__(self, device=None):
        self._L = lib()
        self._C = libcuda()
        if device is None:
            self.handle = self._L.sift3d_make_detector()
        else:
            self.handle = self._L.sift3d_make_detector_on_device(int(device))
        if not self.handle:
            raise Sift3DError("sift3d_make_detector failed (no CUDA device? there is no CPU path)")
        self.ctx = self._L.sift3d_detector_cuda_ctx(self.handle)

    # setters return the C status (0 / -1) like the reference
    def set_peak_thresh(self, v): return int(self._L.sift3d_detector_set_peak_thresh(self.handle, float(v)))
    def set_corner_thresh(self, v): return int(self._L.sift3d_detector_set_corner_thresh(self.handle, float(v)))
    def set_num_kp_levels(self, v): return int(self._L.sift3d_detector_set_num_kp_levels(self.handle, int(v)))
    def set_sigma_n(self, v): return int(self._L.sift3d_detector_set_sigma_n(self.handle, float(v)))
    def set_sigma0(self, v): return int(self._L.sift3d_detector_set_sigma0(self.handle, float(v)))

    def detect(self, image: Image, store: KeypointStore | None = None) -> KeypointStore:
        store = KeypointStore() if store is None else store
        if self._L.sift3d_detect_keypoints(self.handle, image.handle, store.handle):
            raise Sift3DError("sift3d_detect_keypoints failed")
        return store

    def extract(self, kp: KeypointStore, store: DescriptorStore | None = None) -> DescriptorStore:
        store = DescriptorStore() if store is None else store
        if self._L.sift3d_extract_descriptors(self.handle, kp.handle, store.handle):
            raise Sift3DError("sift3d_extract_descriptors failed")
        return store

    # ---- inspection through the inner CUDA ABI (tests / bench) ----
    def level(self, o, s):
        dims = (C.c_int * 3)()
        units = (C.c_double * 3)()
        scale = C.c_double()
        if self._C.sift3d_cuda_level_info(self.ctx, o, s, dims, units, C.byref(scale)):
            raise Sift3DError(f"no level ({o}, {s})")
        out = np.empty((dims[2], dims[1], dims[0]), dtype=np.float32)
        if self._C.sift3d_cuda_fetch_level(self.ctx, o, s, out.ctypes.data_as(C.c_void_p)):
            raise Sift3DError("sift3d_cuda_fetch_level failed")
        return out, tuple(units), scale.value

    def num_octaves(self):
        o = 0
        while self._C.sift3d_cuda_level_info(self.ctx, o, 0, None, None, None) == 0:
            o += 1
        return o

    def dogmax(self, o, s):
        v = C.c_float()
        if self._C.sift3d_cuda_dogmax(self.ctx, o, s, C.byref(v)):
            raise Sift3DError("sift3d_cuda_dogmax failed")
        return v.value

    def candidates(self) -> np.ndarray:
        """All candidates of the last detect in scan order, with orientation results."""
        n = C.c_int(1 << 30)
        buf = np.zeros(0, dtype=CUDA_KEYPOINT_DTYPE)
        # first call with max=0 is not supported; size from a generous fetch loop
        cap = 1 << 16
        while True:
            buf = np.zeros(cap, dtype=CUDA_KEYPOINT_DTYPE)
            got = self._C.sift3d_cuda_fetch_candidates(self.ctx, buf.ctypes.data_as(C.c_void_p), cap)
            if got < 0:
                raise Sift3DError("sift3d_cuda_fetch_candidates failed")
            if got < cap:
                return buf[:got].copy()
            cap *= 4

    def timings(self) -> dict:
        t = CudaTimings()
        self._C.sift3d_cuda_get_timings(self.ctx, C.byref(t))
        return t.as_dict()

    def device(self) -> int:
        return int(self._C.sift3d_cuda_ctx_device(self.ctx))

    def mark(self, slot):
        if self._C.sift3d_cuda_mark(self.ctx, int(slot)):
            raise Sift3DError("sift3d_cuda_mark failed")

    def elapsed_ms(self, a, b) -> float:
        ms = C.c_float()
        if self._C.sift3d_cuda_elapsed_ms(self.ctx, int(a), int(b), C.byref(ms)):
            raise Sift3DError("sift3d_cuda_elapsed_ms failed")
        return ms.value

    def detect_describe_sharded(self, im, devices, sharded_octaves=2, describe=True, kp=None, ds=None):
        """sift3d_detect_describe_sharded: z-slabs of one volume over `devices` (entries may
        repeat: several slabs on one device). Returns (KeypointStore, DescriptorStore | None);
        stores passed in are reused (their page-locked buffers with them)."""
        kp = KeypointStore() if kp is None else kp
        ds = (DescriptorStore() if ds is None else ds) if describe else None
        devs = (C.c_int * len(devices))(*[int(v) for v in devices])
        if self._L.sift3d_detect_describe_sharded(self.handle, im.handle, devs, len(devices),
                                                  int(sharded_octaves), kp.handle,
                                                  ds.handle if ds is not None else None):
            raise Sift3DError("sift3d_detect_describe_sharded failed")
        return kp, ds

    def sharded_timings(self) -> dict:
        out = (C.c_float * 7)()
        if self._L.sift3d_sharded_timings(self.handle, out):
            raise Sift3DError("no sharded run on this detector")
        return dict(zip(("total_ms", "slabs_ms", "tail_ms", "describe_ms", "max_detect_ms",
                         "max_describe_ms", "exchanged_mib"), [float(v) for v in out]))

    def set_exact(self, exact=True):
        """0-ULP blur arithmetic (True) or one FFMA per tap (False, the default unless
        $SIFT3D_CUDA_EXACT is set); see include/sift3d_cuda.h."""
        if self._C.sift3d_cuda_set_exact(self.ctx, int(bool(exact))):
            raise Sift3DError("sift3d_cuda_set_exact failed")

    def exact(self) -> bool:
        return self._C.sift3d_cuda_get_exact(self.ctx) == 1

    def set_profiling(self, on=True):
        self._C.sift3d_cuda_set_profiling(self.ctx, int(bool(on)))

    def blur_ms(self, o, s):
        """(milliseconds, launches) of the blur that produced level (o, s) in the last detect."""
        ms, n = C.c_float(), C.c_int()
        if self._C.sift3d_cuda_blur_ms(self.ctx, o, s, C.byref(ms), C.byref(n)):
            raise Sift3DError("sift3d_cuda_blur_ms failed (profiling enabled?)")
        return ms.value, n.value

    def launch_count(self) -> int:
        return int(self._C.sift3d_cuda_launch_count(self.ctx))

    def blur(self, vol: np.ndarray, units, taps, force_generic=False) -> np.ndarray:
        vol = np.ascontiguousarray(vol, dtype=np.float32)
        out = np.empty_like(vol)
        f = CudaFilter()
        f.width = len(taps)
        for i, t in enumerate(taps):
            f.taps[i] = float(t)
        u = (C.c_double * 3)(*[float(x) for x in units])
        if self._C.sift3d_cuda_blur(self.ctx, vol.ctypes.data_as(C.c_void_p),
                                    out.ctypes.data_as(C.c_void_p), vol.shape[2], vol.shape[1],
                                    vol.shape[0], u, C.byref(f), int(force_generic)):
            raise Sift3DError("sift3d_cuda_blur failed")
        return out

    def close(self):
        if getattr(self, "handle", None):
            self._L.sift3d_free_detector(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def detect_describe_batch(detectors, images, describe=True, kps=None, dss=None):
    """sift3d_detect_describe_batch: volume i on detector i % len(detectors), one host thread
    per detector inside the C library. Returns (keypoint stores, descriptor stores). `kps` /
    `dss`: stores to fill (one per volume; a store may repeat for volumes of the same detector,
    which run in sequence -- it then holds the last of them)."""
    L = lib()
    n = len(images)
    kps = [KeypointStore() for _ in range(n)] if kps is None else kps
    dss = ([DescriptorStore() for _ in range(n)] if dss is None else dss) if describe else []
    vp = C.c_void_p
    a_det = (vp * len(detectors))(*[d.handle for d in detectors])
    a_im = (vp * max(n, 1))(*[im.handle for im in images])
    a_kp = (vp * max(n, 1))(*[k.handle for k in kps])
    a_ds = (vp * max(n, 1))(*[d.handle for d in dss]) if describe else None
    if L.sift3d_detect_describe_batch(a_det, len(detectors), a_im, a_kp, a_ds, n):
        raise Sift3DError("sift3d_detect_describe_batch failed")
    return kps, dss
