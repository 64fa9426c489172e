"""CPU tests (-m "not gpu"): the C-ABI libraries load and export every symbol that
include/*.h declares; host-only logic of the drop-in library; loud failure without a
GPU. No compute entry point is called here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import sift3d_b200 as s3
from sift3d_b200 import api

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# The 27 symbols `nm -D` shows for the reference's libsift3D.so.2
REFERENCE_SYMBOLS = """
sift3d_make_detector sift3d_free_detector sift3d_detector_set_peak_thresh
sift3d_detector_set_corner_thresh sift3d_detector_set_num_kp_levels sift3d_detector_set_sigma_n
sift3d_detector_set_sigma0 sift3d_detect_keypoints sift3d_extract_descriptors
sift3d_make_keypoint_store sift3d_free_keypoint_store sift3d_keypoint_store_to_mat_rm
sift3d_keypoint_store_save sift3d_keypoint_store_sort_by_strength sift3d_make_descriptor_store
sift3d_free_descriptor_store sift3d_descriptor_store_save sift3d_descriptor_store_to_mat_rm
sift3d_make_image sift3d_free_image sift3d_read_image sift3d_image_data sift3d_make_mat_rm
sift3d_free_mat_rm sift3d_mat_rm_data sift3d_mat_rm_dimensions sift3d_mat_rm_type
""".split()


def _declared(header, macro):
    text = open(os.path.join(ROOT, "include", header)).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"^\s*#.*$", "", text, flags=re.M)
    names = re.findall(macro + r"\s+[^;{]*?\b(\w+)\s*\(", text)
    return sorted(set(names))


def test_reference_symbol_list_is_complete():
    assert len(REFERENCE_SYMBOLS) == 27 and len(set(REFERENCE_SYMBOLS)) == 27


def test_host_library_exports_every_declared_symbol():
    L = C.CDLL(api.lib_paths()["cuda"], mode=C.RTLD_GLOBAL)
    H = C.CDLL(api.lib_paths()["host"])
    declared = _declared("sift3d.h", "SIFT3D_EXPORT")
    assert set(REFERENCE_SYMBOLS) <= set(declared)
    for name in declared:
        assert hasattr(H, name), f"libsift3D.so does not export {name}"
    assert L is not None


def test_cuda_library_exports_every_declared_symbol():
    L = C.CDLL(api.lib_paths()["cuda"], mode=C.RTLD_GLOBAL)
    declared = _declared("sift3d_cuda.h", "SIFT3D_CUDA_EXPORT")
    assert len(declared) >= 15
    for name in declared:
        assert hasattr(L, name), f"libsift3d_cuda.so does not export {name}"


def test_compat_headers_exist():
    for h in ("sift.h", "imutil.h", "imtypes.h"):
        assert os.path.exists(os.path.join(ROOT, "include", "sift3d", h))


def test_abi_struct_sizes():
    assert C.sizeof(api.CudaKeypoint) == 88
    assert api.CUDA_KEYPOINT_DTYPE.itemsize == 88
    assert C.sizeof(api.CudaFilter) == 4 + 4 * api.MAX_TAPS


def test_mat_rm_host_logic():
    L = s3.lib()
    m = L.sift3d_make_mat_rm()
    assert m
    assert L.sift3d_mat_rm_type(m) == 1  # SIFT3D_FLOAT (imutil.c:1678)
    cols, rows = C.c_int(-1), C.c_int(-1)
    L.sift3d_mat_rm_dimensions(m, C.byref(cols), C.byref(rows))
    assert (cols.value, rows.value) == (0, 0)
    L.sift3d_mat_rm_dimensions(m, None, None)  # NULL outputs are allowed (imutil.c:1698-1706)
    L.sift3d_free_mat_rm(m)


def test_empty_stores_host_logic(tmp_path):
    kp = s3.KeypointStore()
    assert len(kp) == 0
    kp.sort_by_strength(100)
    assert kp.to_matrix().shape == (0, 3)
    path = tmp_path / "kp.csv"
    assert kp.save(path) == 0 and path.read_text() == ""
    ds = s3.DescriptorStore()
    assert len(ds) == 0
    with pytest.raises(s3.Sift3DError):
        ds.to_matrix()  # sift.c:1691-1695: fails on an empty store
    assert ds.save(tmp_path / "d.csv") == -1


def test_read_image_rejects_unknown_extension():
    L = s3.lib()
    assert not L.sift3d_read_image(b"/nonexistent/file.xyz")


def test_fails_loudly_without_gpu():
    if s3.cuda_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(s3.Sift3DError):
        s3.Detector()
    with pytest.raises(s3.Sift3DError):
        s3.Image(8, 8, 8)  # page-locked allocation needs the CUDA driver
    assert not s3.libcuda().sift3d_cuda_ctx_create(0)


def test_product_does_not_link_or_import_the_oracle():
    """The product must never route through oracle/ (or any CPU fallback)."""
    for root, _, files in os.walk(os.path.join(ROOT, "sift3d_b200")):
        for f in files:
            if f.endswith((".py", ".c", ".cu", ".cuh", ".h")):
                text = open(os.path.join(root, f), errors="ignore").read()
                hit = re.search(r"(from|import)\s+oracle|oracle[/\\.]\w|libsift3d_oracle|refstage|"
                                r"sift3D_ref", text)
                assert hit is None, (os.path.join(root, f), hit.group(0))
    out = os.popen(f"ldd {api.lib_paths()['host']} {api.lib_paths()['cuda']} 2>/dev/null").read()
    assert "oracle" not in out and "sift3D_ref" not in out and "refstage" not in out


def test_synth_shapes():
    v = s3.synth_volume(20, 12, 9, num_blobs=3, seed=1)
    assert v.shape == (9, 12, 20) and v.dtype == np.float32 and np.isfinite(v).all()
    assert np.abs(v).max() > 0


@pytest.mark.parametrize("tiles,nzr,hw,ncta,top", [
    (32, 256, 3, 148, 1), (64, 256, 2, 444, 1), (32, 256, 8, 134, 1), (128, 300, 5, 148, 1),
    (512, 1024, 8, 148, 1), (4, 40, 2, 20, 1), (2, 17, 8, 2, 1), (1, 9, 4, 1, 0), (6, 96, 6, 40, 0),
    (3, 23, 1, 8, 1),
])
def test_fused_blur_chunk_table(tiles, nzr, hw, ncta, top):
    """Host logic of the fused blur's work table (blur.cu: fused_chunk_table): the ranges tile
    the flattened (tile, plane) space in order, no piece of a range inside one tile is shorter
    than hw + 1 outputs (a chunk below the top of the volume must not reach the virtual planes
    of the right-boundary mirror, imutil.c:846-849), and the cost is balanced."""
    L = api.libcuda()
    L.sift3d_cuda_blur_chunk_table.argtypes = [C.c_int] * 5 + [C.POINTER(C.c_int)]
    L.sift3d_cuda_blur_chunk_table.restype = C.c_int
    b = (C.c_int * (ncta + 1))()
    worst = L.sift3d_cuda_blur_chunk_table(ncta, tiles, nzr, hw, top, b)
    b = list(b)
    total = tiles * nzr
    assert worst > 0 and b[0] == 0 and b[-1] == total
    assert all(b[i] <= b[i + 1] for i in range(ncta))
    minlen = min(hw + 1, nzr)
    costs = []
    for i in range(ncta):
        pos, cost = b[i], 0
        while pos < b[i + 1]:
            rem = nzr - pos % nzr
            take = min(rem, b[i + 1] - pos)
            assert take >= minlen, (i, pos, take)
            cost += take + 2 * hw + (hw + 2 if (take == rem and top) else 0)
            pos += take
        costs.append(cost)
    assert max(costs) == worst
    # balance: the most expensive range is within one tile-crossing of the mean
    busy = [c for c in costs if c > 0]
    mean = sum(busy) / len(busy)
    assert worst <= mean + 3 * (2 * hw + minlen) + hw + 2, (worst, mean)
    assert L.sift3d_cuda_blur_chunk_table(0, tiles, nzr, hw, top, b and (C.c_int * 2)()) == -1


def test_committed_blur_traffic_belongs_to_the_current_kernel():
    """bench.py's roofline.traffic comes from an ncu capture committed under profiles/; the file
    names the kernel source it was taken from and that source's hash (bench.py reports null when
    they differ). Changing the kernel without a fresh capture should be noticed here."""
    import hashlib
    import json
    tj = json.load(open(os.path.join(ROOT, "profiles", "r2_blur_traffic.json")))
    src = os.path.join(ROOT, tj["kernel_source"])
    assert hashlib.sha256(open(src, "rb").read()).hexdigest() == tj["kernel_source_sha256"]
    assert 0.5 * tj["algorithmic_bytes_per_blur"] < tj["dram_bytes_per_blur"] < 2 * tj["algorithmic_bytes_per_blur"]
