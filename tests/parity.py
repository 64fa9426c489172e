"""Shared helpers for the parity tests: run the CUDA product and an oracle on the same
seeded volume and compare stage by stage with the tolerances BASELINE.json states.

    pyramid voxels   <= 1e-5 relative to the level's max-abs (expected: bit-identical)
    candidates       identical set and order
    keypoints        >= 99 % matched within 0.5 voxel and 1e-3 scale
    descriptors      cosine >= 0.999 on matched keypoints
"""
from __future__ import annotations

import numpy as np

VOXEL_REL_TOL = 1e-5
KP_MATCH_FRACTION = 0.99
KP_POS_TOL = 0.5
KP_SCALE_TOL = 1e-3
DESC_COS_MIN = 0.999


def level_error(a: np.ndarray, b: np.ndarray):
    """(max abs err relative to max|b|, number of voxels whose bits differ)."""
    assert a.shape == b.shape, (a.shape, b.shape)
    denom = float(np.max(np.abs(b))) or 1.0
    rel = float(np.max(np.abs(a.astype(np.float64) - b.astype(np.float64)))) / denom
    nbits = int(np.count_nonzero(a.view(np.uint32) != b.view(np.uint32)))
    return rel, nbits


def cand_tuple_set(c):
    return {(int(r["o"]), int(r["s"]), int(r["x"]), int(r["y"]), int(r["z"])) for r in c}


def cand_tuple_list(c):
    return [(int(r["o"]), int(r["s"]), int(r["x"]), int(r["y"]), int(r["z"])) for r in c]


def match_keypoints(a, b):
    """Indices (i, j) of keypoints of a matched to b: same octave and level scale within
    1e-3, base-octave position within 0.5 voxel."""
    key_b = {}
    for j, r in enumerate(b):
        key_b.setdefault((int(r["o"]), int(r["s"]), int(r["x"]), int(r["y"]), int(r["z"])), j)
    pairs = []
    for i, r in enumerate(a):
        j = key_b.get((int(r["o"]), int(r["s"]), int(r["x"]), int(r["y"]), int(r["z"])))
        if j is not None and abs(r["sd"] - b[j]["sd"]) <= KP_SCALE_TOL:
            pairs.append((i, j))
    return pairs


def cosine_rows(a: np.ndarray, b: np.ndarray) -> np.ndarray:
    num = np.sum(a.astype(np.float64) * b.astype(np.float64), axis=1)
    den = np.linalg.norm(a.astype(np.float64), axis=1) * np.linalg.norm(b.astype(np.float64), axis=1)
    return num / np.maximum(den, 1e-300)


def run_product(vol, units=(1.0, 1.0, 1.0), params=None, limit=None, exact=None):
    """Full detect(+sort)+extract through the drop-in C API. Returns a dict.
    exact: True = 0-ULP blur arithmetic, False = fast (FFMA), None = the library default."""
    import sift3d_b200 as s3

    det = s3.Detector()
    if exact is not None:
        det.set_exact(exact)
    for k, v in (params or {}).items():
        assert getattr(det, f"set_{k}")(v) == 0
    im = s3.Image.from_array(vol, units)
    kp = det.detect(im)
    cands = det.candidates()
    out = {"det": det, "candidates": cands, "keys_all": kp.records()}
    if limit is not None:
        kp.sort_by_strength(limit)
    out["keys"] = kp.records()
    out["kp_matrix"] = kp.to_matrix()
    if len(kp):
        ds = det.extract(kp)
        out["desc"] = ds.to_matrix()
    else:
        out["desc"] = np.zeros((0, 771), np.float32)
    out["kp_store"] = kp
    return out


def run_reference(vol, units=(1.0, 1.0, 1.0), params=None, limit=None):
    from oracle import refstage

    r = refstage.RefRun(vol, units, params)
    keys_all = r.detect()
    out = {"run": r, "candidates": r.candidates, "keys_all": keys_all}
    if limit is not None:
        r.sort_by_strength(limit)
    out["keys"] = r.keypoints()
    out["kp_matrix"] = r.keypoint_matrix()
    out["desc"] = r.extract() if len(out["keys"]) else np.zeros((0, 771), np.float32)
    return out


def compare_pyramids(det, ref_run, report=None):
    """Every Gaussian level of every octave. Returns worst (rel, nbits)."""
    worst = (0.0, 0)
    noct = ref_run.num_octaves()
    assert det.num_octaves() == noct
    nlev = ref_run.L.refstage_num_levels(ref_run.det, 0)
    for o in range(noct):
        for s in range(-1, nlev - 1):
            a, ua, sa = det.level(o, s)
            b, ub, sb = ref_run.level(o, s)
            assert ua == ub, (o, s, ua, ub)
            assert sa == sb, (o, s, sa, sb)
            rel, nb = level_error(a, b)
            if report is not None:
                report.append((o, s, rel, nb))
            worst = (max(worst[0], rel), max(worst[1], nb))
    return worst


def compare_runs(got, want, exact, r_atol=2e-5):
    """The parity gates of BASELINE.json's north_star between a product run and a reference
    run (dicts of run_product / run_reference). Exact mode additionally demands bit-identical
    pyramids and identical candidate lists; fast mode (FFMA blur) the stated tolerances:
    voxels within 1e-5 of the level's max-abs, >= 99 % of the candidates and keypoints
    matched (unmatched ones are marginal extrema flipped by float reordering), descriptor
    cosine >= 0.999. Returns a dict of the measured figures."""
    report = []
    rel, nbits = compare_pyramids(got["det"], want["run"], report)
    assert rel <= VOXEL_REL_TOL, report
    gc, wc = cand_tuple_list(got["candidates"]), cand_tuple_list(want["candidates"])
    if exact:
        assert nbits == 0, [r for r in report if r[3]]
        assert gc == wc
        np.testing.assert_array_equal(got["candidates"]["strength"], want["candidates"]["strength"])
    else:
        common = len(set(gc) & set(wc))
        assert common >= KP_MATCH_FRACTION * len(wc), (common, len(wc))
        assert common >= KP_MATCH_FRACTION * len(gc), (common, len(gc))
    pairs = match_keypoints(want["keys"], got["keys"])
    assert len(want["keys"]) > 0
    assert len(pairs) >= KP_MATCH_FRACTION * len(want["keys"]), (len(pairs), len(want["keys"]))
    assert len(got["keys"]) <= len(want["keys"]) / KP_MATCH_FRACTION + 1
    iw = [i for i, _ in pairs]
    ig = [j for _, j in pairs]
    np.testing.assert_allclose(got["keys"]["R"][ig], want["keys"]["R"][iw], atol=r_atol)
    cos = cosine_rows(got["desc"][ig, 3:], want["desc"][iw, 3:])
    assert cos.min() >= DESC_COS_MIN, cos.min()
    np.testing.assert_array_equal(got["desc"][ig, :3], want["desc"][iw, :3])
    return {"voxel_rel": rel, "voxels_differing": nbits, "candidates": (len(gc), len(wc)),
            "keypoints": (len(got["keys"]), len(want["keys"]), len(pairs)),
            "min_cosine": float(cos.min())}
