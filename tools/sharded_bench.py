#!/usr/bin/env python3
"""Config 5: one 1024^3 (or given) volume, z-slab sharded over N GPUs of this box
(sift3d_detect_describe_sharded) against the same volume on one GPU. Prints one JSON line.
    python tools/sharded_bench.py [--dim 1024] [--gpus 1,2,4,8] [--octaves 3] [--reps 3] [--check]"""
import argparse, json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import sift3d_b200 as s3
from sift3d_b200.synth import synth_volume

ap = argparse.ArgumentParser()
ap.add_argument("--dim", type=int, default=1024)
ap.add_argument("--nz", type=int, default=0)
ap.add_argument("--gpus", default="")
ap.add_argument("--octaves", type=int, default=3)
ap.add_argument("--reps", type=int, default=3)
ap.add_argument("--check", action="store_true", help="compare with the one-device result (exact mode)")
args = ap.parse_args()
nx = ny = args.dim
nz = args.nz or args.dim
ndev = s3.cuda_device_count()
gpus = [int(g) for g in args.gpus.split(",")] if args.gpus else [n for n in (1, 2, 4, 8) if n <= ndev]
im = s3.Image(nx, ny, nz)
t0 = time.time()
synth_volume(nx, ny, nz, num_blobs=nx * ny * nz // 4096, flags=0, out=im.array)
rep = {"dims": [nx, ny, nz], "synth_s": round(time.time() - t0, 1), "sharded_octaves": args.octaves,
       "mode": "exact" if os.environ.get("SIFT3D_CUDA_EXACT") else "fast", "runs": {}}
ref = None
det = s3.Detector(0)
kp, ds = s3.KeypointStore(), s3.DescriptorStore()
best = None
for _ in range(args.reps):
    t0 = time.perf_counter()
    det.detect(im, kp)
    det.extract(kp, ds)
    dt = time.perf_counter() - t0
    best = dt if best is None else min(best, dt)
rep["runs"]["1_single_device"] = {"e2e_ms": round(best * 1e3, 2), "keypoints": len(kp),
                                  "Gvox_s": round(nx * ny * nz / best / 1e9, 3)}
if args.check:
    ref = (kp.records().copy(), ds.to_matrix().copy())
det.close()
for n in gpus:
    if n < 1 or n > ndev:
        continue
    det = s3.Detector(0)
    best, tm = None, None
    kp, ds = s3.KeypointStore(), s3.DescriptorStore()
    for _ in range(args.reps + 1):  # the first run builds the contexts
        t0 = time.perf_counter()
        det.detect_describe_sharded(im, list(range(n)), args.octaves, kp=kp, ds=ds)
        dt = time.perf_counter() - t0
        if _ > 0 and (best is None or dt < best):
            best, tm = dt, det.sharded_timings()
    r = {"e2e_ms": round(best * 1e3, 2), "keypoints": len(kp), "Gvox_s": round(nx * ny * nz / best / 1e9, 3),
         "timings": {k: round(v, 2) for k, v in tm.items()}}
    if ref is not None:
        keys, desc = kp.records(), ds.to_matrix()
        same = len(keys) == len(ref[0]) and all(np.array_equal(keys[f], ref[0][f]) for f in keys.dtype.names)
        r["keypoints_identical"] = bool(same)
        r["descriptors_identical"] = bool(same and np.array_equal(desc.view(np.uint32), ref[1].view(np.uint32)))
        if same and not r["descriptors_identical"]:
            r["descriptor_max_abs_diff"] = float(np.abs(desc - ref[1]).max())
    rep["runs"][f"{n}_sharded"] = r
    det.close()
print(json.dumps(rep))
