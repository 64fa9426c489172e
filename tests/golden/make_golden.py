#!/usr/bin/env python3
"""Generate the golden vectors under tests/golden/ from the UNMODIFIED reference
(oracle/_ref/libsift3d_refstage.so, built from /root/reference by oracle/Makefile).

The reference ships no tests, fixtures or golden files, so these outputs
of the reference itself, run in the build container, are what pins the oracle
restatement (oracle/sift3d_oracle.c) and, through it, the CUDA path.

    python tests/golden/make_golden.py          # rewrites tests/golden/*.npz

Contents of each pipeline_*.npz: the input volume, units, every Gaussian level of every
octave, the candidate list (between detect_extrema and assign_orientations), the
keypoints (+R), the N x 3 keypoint matrix and the N x 771 descriptor matrix.
kats.npz: Gaussian taps of the default detector, the icosahedron table, blur
impulse/ramp responses for several (half-width, units) pairs and LAPACK eigen results.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from oracle import refstage  # noqa: E402
from sift3d_b200.synth import synth_volume  # noqa: E402

PIPELINE_CASES = {
    # name: (dims (nx, ny, nz), units, blobs, flags, seed, params)
    "iso_40x36x32": ((40, 36, 32), (1.0, 1.0, 1.0), 60, 0, 4242, None),
    "aniso_32x40x24": ((32, 40, 24), (0.75, 0.75, 1.5), 50, 0, 777, None),
    "params_36": ((36, 36, 36), (1.0, 1.0, 1.0), 60, 2, 99,
                  {"peak_thresh": 0.05, "corner_thresh": 0.3, "num_kp_levels": 2,
                   "sigma0": 1.4, "sigma_n": 0.9}),
}


def gauss_taps(sigma):
    hw = max(int(np.ceil(sigma * 3.0)), 1)
    x = (np.arange(2 * hw + 1, dtype=np.float64) - hw) / (sigma + np.finfo(np.float64).eps)
    k = np.exp(-0.5 * x * x).astype(np.float32)
    acc = np.float32(0)
    for v in k:
        acc = np.float32(acc + v)
    return (k / acc).astype(np.float32)


def make_pipeline(name, dims, units, blobs, flags, seed, params):
    vol = synth_volume(*dims, num_blobs=blobs, seed=seed, flags=flags)
    run = refstage.RefRun(vol, units, params)
    keys = run.detect()
    out = {
        "volume": vol, "units": np.array(units, dtype=np.float64),
        "candidates": run.candidates, "keypoints": keys,
        "kp_matrix": run.keypoint_matrix(),
        "num_octaves": np.int32(run.num_octaves()),
        "num_levels": np.int32(run.L.refstage_num_levels(run.det, 0)),
        "params": np.array(repr(params or {})),
    }
    for o in range(run.num_octaves()):
        for s in range(-1, int(out["num_levels"]) - 1):
            arr, u, sc = run.level(o, s)
            out[f"gpyr_o{o}_s{s}"] = arr
            out[f"gpyr_o{o}_s{s}_meta"] = np.array(list(u) + [sc], dtype=np.float64)
    out["descriptors"] = run.extract() if len(keys) else np.zeros((0, 771), np.float32)
    # CLI sequence (cli/kpSift3D.c:116-135): sort by strength, keep 5, describe
    run.sort_by_strength(5)
    out["keypoints_top5"] = run.keypoints()
    out["descriptors_top5"] = run.extract() if len(keys) else np.zeros((0, 771), np.float32)
    print(f"{name}: {len(run.candidates)} candidates, {len(keys)} keypoints, "
          f"{run.num_octaves()} octaves")
    np.savez_compressed(os.path.join(HERE, f"pipeline_{name}.npz"), **out)


def make_kats():
    out = {}
    run = refstage.RefRun(np.zeros((16, 16, 16), np.float32))
    run.detect()
    for i in range(-1, 5):
        taps, sigma = run.filter(i)
        out[f"taps_{i + 1}"] = taps
        out[f"sigma_{i + 1}"] = np.float64(sigma)
    v, idx = run.mesh()
    out["mesh_v"] = v
    out["mesh_idx"] = idx
    rng = np.random.default_rng(2024)
    blur_cases = []
    for ci, (sigma, units) in enumerate([(0.5387, (1, 1, 1)), (0.9733, (2, 2, 2)),
                                          (1.5450, (4, 4, 4)), (2.4525, (1, 1, 1)),
                                          (1.2263, (0.75, 0.75, 1.5)), (2.4525, (2, 2, 2)),
                                          (1.9466, (1.0, 2.0, 0.5))]):
        taps = gauss_taps(sigma)
        dims = (21, 19, 18)  # (nz, ny, nx)
        impulse = np.zeros(dims, np.float32)
        impulse[10, 9, 9] = 1.0
        impulse[0, 0, 0] = 2.0
        impulse[-1, -1, -1] = -3.0
        ramp = (np.arange(dims[2], dtype=np.float32)[None, None, :] * 0.25 +
                np.arange(dims[1], dtype=np.float32)[None, :, None] * -0.5 +
                np.arange(dims[0], dtype=np.float32)[:, None, None] * 0.125)
        ramp = np.ascontiguousarray(np.broadcast_to(ramp, dims), dtype=np.float32)
        noise = rng.standard_normal(dims).astype(np.float32)
        for nm, src in (("impulse", impulse), ("ramp", ramp), ("noise", noise)):
            out[f"blur{ci}_{nm}_in"] = src
            out[f"blur{ci}_{nm}_out"] = refstage.ref_blur(src, units, taps)
        out[f"blur{ci}_taps"] = taps
        out[f"blur{ci}_units"] = np.array(units, dtype=np.float64)
        blur_cases.append(ci)
    out["blur_cases"] = np.array(blur_cases, dtype=np.int32)
    mats = []
    for _ in range(16):
        g = rng.standard_normal((3, 3))
        mats.append(g @ g.T * rng.uniform(1e-6, 1e3))
    mats = np.array(mats)
    out["eig_A"] = mats
    out["eig_w"] = np.array([refstage.ref_eigen3(m)[0] for m in mats])
    out["eig_Q"] = np.array([refstage.ref_eigen3(m)[1] for m in mats])
    np.savez_compressed(os.path.join(HERE, "kats.npz"), **out)
    print("kats.npz written")


if __name__ == "__main__":
    assert refstage.available(), "build oracle/_ref first: make -C oracle ref"
    for name, cfg in PIPELINE_CASES.items():
        make_pipeline(name, *cfg)
    make_kats()
