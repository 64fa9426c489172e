"""Batch sharding of independent volumes across the GPUs of one box (BASELINE.json
configs[3]): one process / detector / stream / arena per GPU, volumes dealt
round-robin, NO data-path collective. torch.distributed is used only to agree on the
assignment and to collect the per-volume summaries on rank 0.

The reference has no counterpart (it is single-process; distinct detectors are
independent because the library has no mutable globals, sift.c:1841 / imtypes_private.h)."""
from __future__ import annotations

from typing import Callable, Dict, List, Sequence


def shard_indices(num_volumes: int, world_size: int, rank: int) -> List[int]:
    """Volumes owned by `rank`: i with i % world_size == rank (round-robin keeps the
    per-GPU load even when the batch is not a multiple of the world size)."""
    if world_size < 1 or not 0 <= rank < world_size:
        raise ValueError(f"bad rank {rank} / world size {world_size}")
    return list(range(rank, max(num_volumes, 0), world_size))


def run_shard(indices: Sequence[int], process: Callable[[int], dict]) -> Dict[int, dict]:
    """Run `process(volume_index) -> summary dict` for every owned volume, in order."""
    return {i: process(i) for i in indices}


def gather_summaries(local: Dict[int, dict], dist=None) -> List[dict] | None:
    """Collect the per-volume summaries of all ranks on rank 0, ordered by volume index.
    `dist` is torch.distributed (already initialised) or None for a single process.
    Returns the list on rank 0 and None elsewhere."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return [local[i] for i in sorted(local)]
    world = dist.get_world_size()
    parts = [None] * world
    dist.all_gather_object(parts, local)
    if dist.get_rank() != 0:
        return None
    merged: Dict[int, dict] = {}
    for p in parts:
        for k, v in p.items():
            if k in merged:
                raise RuntimeError(f"volume {k} was processed by two ranks")
            merged[k] = v
    return [merged[i] for i in sorted(merged)]


def detect_and_describe(det, volume, units=(1.0, 1.0, 1.0)) -> dict:
    """One volume through the public API on this rank's detector; summary only (counts and
    a checksum of the descriptor matrix), the stores stay with the caller if needed."""
    import numpy as np

    from .api import Image

    kp = det.detect(Image.from_array(volume, units))
    out = {"keypoints": len(kp), "desc_checksum": 0.0}
    if len(kp):
        m = det.extract(kp).to_matrix()
        out["desc_checksum"] = float(np.float64(m[:, 3:]).sum())
    return out
