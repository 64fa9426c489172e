"""CPU tests (-m "not gpu"): the oracle restatement (oracle/sift3d_oracle.c) against the
golden vectors generated from the unmodified reference (tests/golden/make_golden.py) and,
when oracle/_ref is present, against the reference live."""
import glob
import os

import numpy as np
import pytest

import parity
from oracle import port, refstage
from sift3d_b200.synth import synth_volume

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
PIPELINES = sorted(glob.glob(os.path.join(GOLDEN, "pipeline_*.npz")))


@pytest.fixture(scope="module", autouse=True)
def _need_port():
    assert port.available(), "oracle/libsift3d_oracle.so missing: run `make -C oracle`"


def _params(g):
    return eval(str(g["params"]))  # written by make_golden.py via repr(dict)


@pytest.mark.parametrize("path", PIPELINES, ids=[os.path.basename(p)[9:-4] for p in PIPELINES])
def test_port_matches_golden_pipeline(path):
    g = np.load(path)
    run = port.PortRun(g["volume"], tuple(g["units"]), _params(g))
    keys = run.detect()
    assert run.num_octaves() == int(g["num_octaves"])
    assert run.num_levels() == int(g["num_levels"])
    for o in range(run.num_octaves()):
        for s in range(-1, run.num_levels() - 1):
            arr, units, scale = run.level(o, s)
            want = g[f"gpyr_o{o}_s{s}"]
            meta = g[f"gpyr_o{o}_s{s}_meta"]
            assert tuple(meta[:3]) == units and meta[3] == scale
            rel, nbits = parity.level_error(arr, want)
            assert nbits == 0, (o, s, rel, nbits)
    assert parity.cand_tuple_list(run.candidates) == parity.cand_tuple_list(g["candidates"])
    np.testing.assert_array_equal(run.candidates["strength"], g["candidates"]["strength"])
    np.testing.assert_array_equal(run.candidates["sd"], g["candidates"]["sd"])
    want = g["keypoints"]
    assert len(keys) == len(want)
    for f in ("o", "s", "x", "y", "z", "sd", "strength"):
        np.testing.assert_array_equal(keys[f], want[f])
    # Jacobi vs LAPACK: same eigenvectors to double precision -> same floats up to 1 ulp
    np.testing.assert_allclose(keys["R"], want["R"], atol=1e-6)
    np.testing.assert_array_equal(run.keypoint_matrix(), g["kp_matrix"])
    desc = run.extract()
    np.testing.assert_array_equal(desc[:, :3], g["descriptors"][:, :3])
    np.testing.assert_allclose(desc[:, 3:], g["descriptors"][:, 3:], atol=2e-6)
    assert parity.cosine_rows(desc[:, 3:], g["descriptors"][:, 3:]).min() > 1 - 1e-9
    # CLI sequence: sort by (misaligned) strength, keep 5
    run.sort_by_strength(5)
    top = run.keypoints()
    for f in ("o", "s", "x", "y", "z", "strength"):
        np.testing.assert_array_equal(top[f], g["keypoints_top5"][f])
    np.testing.assert_allclose(run.extract()[:, 3:], g["descriptors_top5"][:, 3:], atol=2e-6)


def test_port_matches_golden_kats():
    k = np.load(os.path.join(GOLDEN, "kats.npz"))
    run = port.PortRun(np.zeros((16, 16, 16), np.float32))
    run.detect()
    for i in range(6):
        np.testing.assert_array_equal(run.filter(i - 1), k[f"taps_{i}"])
    assert [len(k[f"taps_{i}"]) // 2 for i in range(6)] == [2, 3, 4, 5, 6, 8]
    v, idx = port.port_mesh()
    np.testing.assert_array_equal(v, k["mesh_v"])
    np.testing.assert_array_equal(idx, k["mesh_idx"])
    for ci in k["blur_cases"]:
        for nm in ("impulse", "ramp", "noise"):
            got = port.port_blur(k[f"blur{ci}_{nm}_in"], k[f"blur{ci}_units"], k[f"blur{ci}_taps"])
            rel, nbits = parity.level_error(got, k[f"blur{ci}_{nm}_out"])
            assert nbits == 0, (int(ci), nm, rel, nbits)
    for A, w, Q in zip(k["eig_A"], k["eig_w"], k["eig_Q"]):
        gw, gQ = port.port_eigen3(A)
        np.testing.assert_allclose(gw, w, rtol=1e-12, atol=1e-14 * np.abs(w).max())
        for c in range(3):  # eigenvector sign is arbitrary
            dot = abs(float(np.dot(gQ[:, c], Q[:, c])))
            assert abs(dot - 1) < 1e-9


def test_mesh_swap_quirk():
    """tri.v = (V[idx1], V[idx0], V[idx2]) for all 20 faces."""
    v, idx = port.port_mesh()
    gr = np.float32(1.6180339887)
    verts = np.array([[0, 1, gr], [0, -1, gr], [0, 1, -gr], [0, -1, -gr], [1, gr, 0], [-1, gr, 0],
                      [1, -gr, 0], [-1, -gr, 0], [gr, 0, 1], [-gr, 0, 1], [gr, 0, -1],
                      [-gr, 0, -1]], dtype=np.float64)
    verts /= np.linalg.norm(verts, axis=1, keepdims=True)
    for f in range(20):
        np.testing.assert_allclose(v[f, 0], verts[idx[f, 1]], atol=1e-6)
        np.testing.assert_allclose(v[f, 1], verts[idx[f, 0]], atol=1e-6)
        np.testing.assert_allclose(v[f, 2], verts[idx[f, 2]], atol=1e-6)


@pytest.mark.skipif(not refstage.available(), reason="oracle/_ref not built")
@pytest.mark.parametrize("dims,units,blobs,flags", [
    ((48, 48, 48), (1.0, 1.0, 1.0), 40, 0),
    ((37, 50, 29), (1.0, 1.0, 1.0), 30, 2),
    ((40, 40, 32), (0.75, 0.75, 1.5), 30, 0),
    ((32, 32, 32), (2.0, 1.0, 0.5), 20, 0),
])
def test_port_matches_reference_live(dims, units, blobs, flags):
    vol = synth_volume(*dims, num_blobs=blobs, seed=31337 + blobs, flags=flags)
    r = refstage.RefRun(vol, units)
    rk = r.detect()
    p = port.PortRun(vol, units)
    pk = p.detect()
    for o in range(r.num_octaves()):
        for s in range(-1, 5):
            a, ua, sa = p.level(o, s)
            b, ub, sb = r.level(o, s)
            assert ua == ub and sa == sb
            assert parity.level_error(a, b)[1] == 0, (o, s)
    assert parity.cand_tuple_list(p.candidates) == parity.cand_tuple_list(r.candidates)
    np.testing.assert_array_equal(p.candidates["strength"], r.candidates["strength"])
    assert len(pk) == len(rk)
    for f in ("o", "s", "x", "y", "z", "sd", "strength"):
        np.testing.assert_array_equal(pk[f], rk[f])
    np.testing.assert_allclose(pk["R"], rk["R"], atol=1e-6)
    if len(rk):
        np.testing.assert_allclose(p.extract(), r.extract(), atol=2e-6)


@pytest.mark.skipif(not refstage.available(), reason="oracle/_ref not built")
def test_reference_stale_units_quirk():
    """sift.c:651-656: a second image with the same dims but other units re-uses the
    first image's units for octaves >= 1. The restatement mirrors that."""
    v1 = synth_volume(32, 32, 32, num_blobs=20, seed=5)
    v2 = synth_volume(32, 32, 32, num_blobs=20, seed=6)
    r = refstage.RefRun(v1, (1.0, 1.0, 1.0))
    r.detect()
    L = r.L
    np.ctypeslib.as_array(L.sift3d_image_data(r.im), shape=(v2.size,))[:] = v2.ravel()
    L.refstage_image_set_units(r.im, 2.0, 2.0, 2.0)
    r.detect()
    p = port.PortRun(v1, (1.0, 1.0, 1.0))
    p.detect()
    p.vol = v2
    p.units = (2.0, 2.0, 2.0)
    p.detect()
    for o in range(r.num_octaves()):
        for s in range(-1, 5):
            a, ua, _ = p.level(o, s)
            b, ub, _ = r.level(o, s)
            assert ua == ub, (o, s, ua, ub)
            assert parity.level_error(a, b)[1] == 0, (o, s)
    assert r.level(0, 0)[1] == (2.0, 2.0, 2.0) and r.level(1, 0)[1] == (2.0, 2.0, 2.0)


def test_synth_is_deterministic():
    g = np.load(PIPELINES[0])
    name = os.path.basename(PIPELINES[0])
    assert name == "pipeline_aniso_32x40x24.npz"
    vol = synth_volume(32, 40, 24, num_blobs=50, seed=777, flags=0)
    np.testing.assert_array_equal(vol, g["volume"])
