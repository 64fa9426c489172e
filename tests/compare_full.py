#!/usr/bin/env python3
"""Full-size parity report: run a named workload (sift3d_b200.synth.WORKLOADS) or explicit
dims through the CUDA product and through the CPU oracle (oracle/_ref) on this box and print
the stage-by-stage comparison as one JSON line. Test infrastructure (uses oracle/).

    python tests/compare_full.py C2_mri_256 [--no-pyramid] [--limit 100]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
sys.path.insert(0, os.path.dirname(HERE))

import parity  # noqa: E402
from sift3d_b200.synth import WORKLOADS, workload  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("name", choices=sorted(WORKLOADS))
    ap.add_argument("--no-pyramid", action="store_true")
    ap.add_argument("--limit", type=int, default=None)
    ap.add_argument("--port", action="store_true", help="use the restatement instead of _ref")
    args = ap.parse_args()

    vol, units = workload(args.name)
    t0 = time.time()
    got = parity.run_product(vol, units, limit=args.limit)
    t_gpu = time.time() - t0
    t0 = time.time()
    if args.port:
        from oracle import port
        run = port.PortRun(vol, units)
        keys_all = run.detect()
        want = {"run": run, "candidates": run.candidates, "keys_all": keys_all}
        if args.limit is not None:
            run.sort_by_strength(args.limit)
        want["keys"] = run.keypoints()
        want["desc"] = run.extract()
    else:
        want = parity.run_reference(vol, units, limit=args.limit)
    t_cpu = time.time() - t0

    rep = {"workload": args.name, "dims": list(vol.shape[::-1]), "units": list(units),
           "gpu_wall_s_incl_setup": round(t_gpu, 3), "cpu_wall_s": round(t_cpu, 3),
           "candidates": [len(got["candidates"]), len(want["candidates"])],
           "keypoints_before_limit": [len(got["keys_all"]), len(want["keys_all"])],
           "keypoints": [len(got["keys"]), len(want["keys"])]}
    if not args.no_pyramid:
        levels = []
        worst = parity.compare_pyramids(got["det"], want["run"], levels) if not args.port else None
        if args.port:
            worst = [0.0, 0]
            for o in range(want["run"].num_octaves()):
                for s in range(-1, 5):
                    a, _, _ = got["det"].level(o, s)
                    b, _, _ = want["run"].level(o, s)
                    rel, nb = parity.level_error(a, b)
                    worst = [max(worst[0], rel), max(worst[1], nb)]
        rep["pyramid_max_rel_err"] = worst[0]
        rep["pyramid_voxels_with_different_bits"] = worst[1]
    rep["candidates_identical"] = (parity.cand_tuple_list(got["candidates"]) ==
                                   parity.cand_tuple_list(want["candidates"]))
    if rep["candidates_identical"]:
        rep["candidate_strength_identical"] = bool(np.array_equal(
            got["candidates"]["strength"], want["candidates"]["strength"]))
    pairs = parity.match_keypoints(want["keys"], got["keys"])
    rep["keypoints_matched"] = len(pairs)
    rep["keypoints_matched_fraction"] = len(pairs) / max(1, len(want["keys"]))
    if pairs:
        iw = [i for i, _ in pairs]
        ig = [j for _, j in pairs]
        rep["R_max_abs_diff"] = float(np.abs(got["keys"]["R"][ig] - want["keys"]["R"][iw]).max())
        cos = parity.cosine_rows(got["desc"][ig, 3:], want["desc"][iw, 3:])
        rep["desc_cos_min"] = float(cos.min())
        rep["desc_cos_mean"] = float(cos.mean())
        rep["desc_max_abs_diff"] = float(np.abs(got["desc"][ig, 3:] - want["desc"][iw, 3:]).max())
    rep["gpu_stage_ms"] = got["det"].timings()
    print(json.dumps(rep))


if __name__ == "__main__":
    main()
