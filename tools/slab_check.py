#!/usr/bin/env python
"""z-slab detect + describe of one volume over the ranks of a process group, checked on
rank 0 against the single-device detector on the whole volume: candidates, orientation
matrices, keypoints and descriptors must be identical bit for bit.

  torchrun --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
      tools/slab_check.py --dims 160 160 512 --octaves 2            (NCCL, one GPU per rank)

tests/test_gpu_slab.py calls `run_check` from spawned processes that share one GPU (gloo)."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def single_device(vol, units, params, device):
    from sift3d_b200 import api

    det = api.Detector(device)
    for k, v in (params or {}).items():
        if getattr(det, "set_" + k)(v):
            raise RuntimeError(f"bad parameter {k}")
    img = api.Image.from_array(vol, units)
    t0 = time.perf_counter()
    kp = det.detect(img)
    cands = det.candidates()
    desc = det.extract(kp).to_matrix() if len(kp) else np.zeros((0, 771))
    ms = (time.perf_counter() - t0) * 1e3
    return cands, kp.records(), np.asarray(desc), ms


def compare(result, cands, kps, desc):
    """Bitwise comparison; returns a dict of mismatch counts (all zero = identical)."""
    out = {"candidates": [len(result["candidates"]), len(cands)],
           "keypoints": [len(result["keypoints"]), len(kps)]}
    bad = {}
    if len(result["candidates"]) != len(cands):
        bad["num_candidates"] = 1
    else:
        for f in ("o", "s", "x", "y", "z", "sd", "strength", "valid"):
            bad["cand_" + f] = int(np.count_nonzero(result["candidates"][f] != cands[f]))
        v = cands["valid"] != 0
        bad["cand_R"] = int(np.count_nonzero(
            result["candidates"]["R"][v].view(np.uint32) != cands["R"][v].view(np.uint32)))
    k = result["keypoints"]
    if len(k) != len(kps):
        bad["num_keypoints"] = 1
    elif len(k):
        for f in ("x", "y", "z", "sd", "o", "s", "strength"):
            bad["kp_" + f] = int(np.count_nonzero(k[f] != kps[f]))
        bad["kp_R"] = int(np.count_nonzero(k["R"].view(np.uint32) != kps["R"].view(np.uint32)))
        d = result["descriptors"]
        if desc.shape[0] != d.shape[0]:
            bad["num_descriptors"] = 1
        else:
            theirs = np.ascontiguousarray(desc[:, 3:], dtype=np.float32)
            # Descriptor bins are fp32 sums whose order follows the window's row box; the
            # reference's float window bounds (sift.c:91-99) can move by one empty plane when
            # the volume is translated, so a handful of bins may differ in the last bit.
            out["descriptor_bits_differing"] = int(np.count_nonzero(d.view(np.uint32) != theirs.view(np.uint32)))
            out["descriptor_max_abs_diff"] = float(np.max(np.abs(d - theirs))) if d.size else 0.0
            bad["descriptor_values"] = int(np.count_nonzero(np.abs(d - theirs) > 1e-6))
            scale = np.ldexp(1.0, k["o"].astype(np.int64))
            xyz = np.stack([k["x"] * scale, k["y"] * scale, k["z"] * scale], axis=1)  # sift.c:1530-1533
            bad["descriptor_xyz"] = int(np.count_nonzero(xyz.astype(desc.dtype) != desc[:, :3]))
    out["mismatches"] = bad
    out["identical"] = not any(bad.values())
    return out


def run_check(dims, units, sharded_octaves, device, params=None, seed=1234, halo=None, reps=1):
    """Call inside an initialised process group. Returns the report on rank 0, None elsewhere."""
    import torch
    import torch.distributed as dist

    from sift3d_b200 import slab, synth

    rank = dist.get_rank()
    torch.cuda.set_device(device)
    nx, ny, nz = dims
    vol = synth.synth_volume(nx, ny, nz, seed=seed, flags=3)  # every rank builds the same volume
    sd = slab.SlabDetector(dims, units, sharded_octaves=sharded_octaves, halo=halo, device=device,
                           **(params or {}))
    mine = np.ascontiguousarray(vol[sd.slab.a0:sd.slab.a1])
    ms = []
    for _ in range(reps):
        dist.barrier()
        t0 = time.perf_counter()
        res = sd.detect_and_describe(mine)
        dist.barrier()
        ms.append((time.perf_counter() - t0) * 1e3)
    info = {"world": dist.get_world_size(), "backend": dist.get_backend(), "dims": list(dims),
            "sharded_octaves": sd.K, "tail_octaves": sd.tail_octaves, "halo": sd.halo,
            "layout": [[s.g0, s.g1, s.a0, s.a1] for s in sd.layout],
            "halo_bytes_rank0": sd.exchanged_bytes // reps, "slab_wall_ms": ms}
    sd.close()
    if rank != 0:
        return None
    cands, kps, desc, one_ms = single_device(vol, units, params, device)
    info.update(compare(res, cands, kps, desc))
    info["single_device_wall_ms"] = one_ms
    return info


def main():
    import torch
    import torch.distributed as dist

    ap = argparse.ArgumentParser()
    ap.add_argument("--dims", type=int, nargs=3, default=[160, 160, 512])
    ap.add_argument("--units", type=float, nargs=3, default=[1.0, 1.0, 1.0])
    ap.add_argument("--octaves", type=int, default=2, help="octaves run on slabs")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    try:
        rep = run_check(tuple(a.dims), tuple(a.units), a.octaves, local, reps=a.reps)
        if rep is not None:
            line = json.dumps(rep)
            print(line)
            if a.out:
                with open(a.out, "w") as f:
                    f.write(line + "\n")
            if not rep["identical"]:
                sys.exit(1)
    finally:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
