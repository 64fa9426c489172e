"""NIfTI-1 / Analyze reader and writer behind sift3d_read_image.

The reference reads volumes through nifticlib (nifti.c:52-167), which is absent here, so there is
no reference output to pin against: the checker is a numpy restatement of what the wrapper takes
from a nifti_image -- value = (float) ((double) raw * slope + inter) with slope 0 -> 1, units =
pixdim[1..3] (0 / non-finite -> 1), dimensionality = last dim > 1 (> 4 is an error), 4-D = channels.
Runs without a GPU (sift3d_nifti_read_raw uses malloc)."""
import ctypes as C
import gzip
import os
import struct

import numpy as np
import pytest

from sift3d_b200 import api

DTYPES = {2: "u1", 256: "i1", 4: "i2", 512: "u2", 8: "i4", 768: "u4", 1024: "i8", 1280: "u8",
          16: "f4", 64: "f8"}


def _lib():
    L = api._load("host")
    L.sift3d_nifti_read_raw.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_double),
                                        C.POINTER(C.POINTER(C.c_float))]
    L.sift3d_nifti_write_raw.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.POINTER(C.c_double),
                                         C.POINTER(C.c_float)]
    L.sift3d_nifti_free.argtypes = [C.POINTER(C.c_float)]
    L.sift3d_nifti_free.restype = None
    return L


def read_raw(path):
    L = _lib()
    dims = (C.c_int * 4)()
    units = (C.c_double * 3)()
    data = C.POINTER(C.c_float)()
    rc = L.sift3d_nifti_read_raw(os.fsencode(path), dims, units, C.byref(data))
    if rc:
        return None
    nx, ny, nz, nc = dims
    arr = np.ctypeslib.as_array(data, shape=(nz, ny, nx, nc)).copy()
    L.sift3d_nifti_free(data)
    return arr, tuple(units)


def header(dims, code, pixdim, slope, inter, magic, vox_offset, big=False):
    e = ">" if big else "<"
    h = bytearray(348)
    struct.pack_into(e + "i", h, 0, 348)
    dim = [len(dims)] + list(dims) + [1] * (7 - len(dims))
    struct.pack_into(e + "8h", h, 40, *dim)
    struct.pack_into(e + "h", h, 70, code)
    struct.pack_into(e + "h", h, 72, 8 * np.dtype(DTYPES.get(code, "u1")).itemsize)
    pd = [1.0] + list(pixdim) + [1.0] * (7 - len(pixdim))
    struct.pack_into(e + "8f", h, 76, *pd)
    struct.pack_into(e + "f", h, 108, float(vox_offset))
    struct.pack_into(e + "f", h, 112, slope)
    struct.pack_into(e + "f", h, 116, inter)
    h[344:348] = magic
    return bytes(h)


def write_file(path, payload):
    opener = gzip.open if str(path).endswith(".gz") else open
    with opener(path, "wb") as f:
        f.write(payload)


def expected(raw, slope, inter):
    """raw: array shaped (nc, nz, ny, nx) in file order -> (nz, ny, nx, nc) float32."""
    s = 1.0 if slope == 0.0 else float(np.float32(slope))
    v = (raw.astype(np.float64) * s + float(np.float32(inter))).astype(np.float32)
    return np.ascontiguousarray(np.moveaxis(v, 0, -1))


def sample(rng, code, shape):
    dt = np.dtype(DTYPES[code])
    if dt.kind == "f":
        return rng.standard_normal(shape).astype(dt) * 100
    info = np.iinfo(dt)
    return rng.integers(info.min, info.max, size=shape, dtype=dt, endpoint=True)


@pytest.mark.parametrize("code", sorted(DTYPES))
@pytest.mark.parametrize("big", [False, True])
def test_single_file_all_datatypes_both_byte_orders(tmp_path, code, big):
    rng = np.random.default_rng(code + big)
    nx, ny, nz = 7, 5, 4
    raw = sample(rng, code, (1, nz, ny, nx))
    slope, inter = 0.5, -3.25
    payload = header((nx, ny, nz), code, (0.7, 0.8, 2.5), slope, inter, b"n+1\0", 352, big) + b"\0" * 4
    payload += raw.astype(raw.dtype.newbyteorder(">" if big else "<")).tobytes()
    path = tmp_path / ("a.nii.gz" if code % 3 else "a.nii")
    write_file(path, payload)
    got, units = read_raw(path)
    want = expected(raw, slope, inter)
    assert got.shape == (nz, ny, nx, 1)
    assert got.tobytes() == want.tobytes()
    assert units == tuple(float(np.float32(u)) for u in (0.7, 0.8, 2.5))


def test_zero_slope_is_ignored_and_bad_pixdim_becomes_one(tmp_path):
    rng = np.random.default_rng(1)
    raw = sample(rng, 4, (1, 3, 4, 5))
    payload = header((5, 4, 3), 4, (0.0, float("nan"), 1.5), 0.0, 7.0, b"n+1\0", 352) + b"\0" * 4 + raw.tobytes()
    write_file(tmp_path / "z.nii", payload)
    got, units = read_raw(tmp_path / "z.nii")
    assert got.tobytes() == expected(raw, 0.0, 7.0).tobytes()
    assert units == (1.0, 1.0, 1.5)


def test_extension_bytes_before_the_data_are_skipped(tmp_path):
    rng = np.random.default_rng(2)
    raw = sample(rng, 16, (1, 3, 3, 3))
    ext = b"\1\0\0\0" + struct.pack("<ii", 32, 6) + b"x" * 24   # one 32-byte extension
    payload = header((3, 3, 3), 16, (1, 1, 1), 1.0, 0.0, b"n+1\0", 352 + 32) + ext + raw.tobytes()
    write_file(tmp_path / "e.nii", payload)
    got, _ = read_raw(tmp_path / "e.nii")
    assert got.tobytes() == expected(raw, 1.0, 0.0).tobytes()


def test_four_dimensional_file_becomes_channels(tmp_path):
    rng = np.random.default_rng(3)
    nx, ny, nz, nt = 4, 3, 5, 2
    raw = sample(rng, 512, (nt, nz, ny, nx))
    payload = header((nx, ny, nz, nt), 512, (1, 1, 1, 0), 2.0, 1.0, b"n+1\0", 352) + b"\0" * 4 + raw.tobytes()
    write_file(tmp_path / "t.nii.gz", payload)
    got, _ = read_raw(tmp_path / "t.nii.gz")
    assert got.shape == (nz, ny, nx, nt)
    assert got.tobytes() == expected(raw, 2.0, 1.0).tobytes()
    # trailing singleton dimensions do not count (nifti.c:66-71)
    payload = header((nx, ny, nz, 1, 1), 512, (1, 1, 1, 1, 1), 1.0, 0.0, b"n+1\0", 352) + b"\0" * 4 + raw[:1].tobytes()
    write_file(tmp_path / "s.nii", payload)
    got, _ = read_raw(tmp_path / "s.nii")
    assert got.shape == (nz, ny, nx, 1)


@pytest.mark.parametrize("magic,gz", [(b"ni1\0", False), (b"ni1\0", True), (b"\0\0\0\0", False)])
def test_header_image_pairs_nifti_and_analyze(tmp_path, magic, gz):
    rng = np.random.default_rng(4)
    raw = sample(rng, 8, (1, 4, 3, 6))
    suffix = ".gz" if gz else ""
    write_file(tmp_path / ("p.hdr" + suffix), header((6, 3, 4), 8, (2, 2, 2), 3.0, 1.0, magic, 0))
    write_file(tmp_path / ("p.img" + suffix), raw.tobytes())
    analyze = magic[0] == 0
    want = expected(raw, 0.0 if analyze else 3.0, 0.0 if analyze else 1.0)  # Analyze has no scaling
    for name in ("p.img", "p.hdr"):
        got, units = read_raw(tmp_path / (name + suffix))
        assert got.tobytes() == want.tobytes()
        assert units == (2.0, 2.0, 2.0)


def test_rejected_files(tmp_path, capfd):
    raw = np.zeros((1, 2, 2, 2), np.float32)
    ok = header((2, 2, 2), 16, (1, 1, 1), 1, 0, b"n+1\0", 352) + b"\0" * 4 + raw.tobytes()
    cases = {
        "five_d.nii": header((2, 2, 2, 1, 2), 16, (1,) * 5, 1, 0, b"n+1\0", 352) + b"\0" * 4 + raw.tobytes() * 2,
        "complex.nii": header((2, 2, 2), 32, (1, 1, 1), 1, 0, b"n+1\0", 352) + b"\0" * 4 + raw.tobytes() * 2,
        "short.nii": ok[:-5],
        "garbage.nii": b"not a header" * 40,
        "hdr_only.nii": header((2, 2, 2), 16, (1, 1, 1), 1, 0, b"ni1\0", 0),
    }
    for name, payload in cases.items():
        write_file(tmp_path / name, payload)
        assert read_raw(tmp_path / name) is None, name
    assert read_raw(tmp_path / "missing.nii") is None
    err = capfd.readouterr().err
    assert "unsupported dimensionality 5" in err and "unsupported datatype 32" in err


@pytest.mark.parametrize("name,nc", [("w.nii", 1), ("w.nii.gz", 1), ("w4.nii.gz", 3)])
def test_writer_round_trip_and_header(tmp_path, name, nc):
    L = _lib()
    rng = np.random.default_rng(5)
    nx, ny, nz = 9, 6, 5
    vol = rng.standard_normal((nz, ny, nx, nc)).astype(np.float32)
    dims = (C.c_int * 4)(nx, ny, nz, nc)
    units = (C.c_double * 3)(0.75, 0.75, 1.5)
    path = tmp_path / name
    assert L.sift3d_nifti_write_raw(os.fsencode(path), dims, units, vol.ctypes.data_as(C.POINTER(C.c_float))) == 0
    got, u = read_raw(path)
    assert got.tobytes() == vol.tobytes() and u == (0.75, 0.75, 1.5)
    opener = gzip.open if name.endswith(".gz") else open
    with opener(path, "rb") as f:
        h = f.read(352)
    assert struct.unpack_from("<i", h, 0)[0] == 348 and h[344:348] == b"n+1\0"
    assert struct.unpack_from("<8h", h, 40)[:5] == ((4 if nc > 1 else 3), nx, ny, nz, nc)
    assert struct.unpack_from("<hh", h, 70) == (16, 32)
    assert struct.unpack_from("<ff", h, 112) == (1.0, 0.0)      # write_nii: no scaling (nifti.c:185-186)
    assert L.sift3d_nifti_write_raw(os.fsencode(tmp_path / "w.img"), dims, units,
                                    vol.ctypes.data_as(C.POINTER(C.c_float))) != 0
