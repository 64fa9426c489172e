"""sift3d_b200 -- Python (ctypes) mirror of the drop-in C API, for tests and bench.py.

The product is the pair of shared libraries under ``sift3d_b200/lib`` (built by the
top-level Makefile): ``libsift3D.so`` (host C, the reference's 27 public functions,
``include/sift3d.h``) on top of ``libsift3d_cuda.so`` (hand-written sm_100a kernels,
``include/sift3d_cuda.h``). This module only binds them; it contains no compute and no
fallback: if the libraries are missing, or no CUDA device is present when a detector is
created, it raises.
"""
from .api import (  # noqa: F401
    CudaKeypoint,
    CudaTimings,
    DescriptorStore,
    Detector,
    Image,
    KeypointStore,
    Sift3DError,
    cuda_device_count,
    detect_describe_batch,
    lib,
    lib_paths,
    libcuda,
)
from .synth import synth_volume, workload  # noqa: F401

__all__ = [
    "CudaKeypoint", "CudaTimings", "DescriptorStore", "Detector", "Image", "KeypointStore",
    "Sift3DError", "cuda_device_count", "detect_describe_batch", "lib", "libcuda", "lib_paths", "synth_volume", "workload",
]
