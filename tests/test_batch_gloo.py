"""Multi-process host logic of the batch path (configs[3]) on CPU: world_size-2 gloo.
Every volume is owned by exactly one rank and rank 0 receives all summaries in order;
no GPU and no oracle involved."""
import os
import socket

import pytest

from sift3d_b200 import batch


def test_shard_indices_cover_and_balance():
    for n in (0, 1, 7, 64):
        for world in (1, 2, 3, 8):
            owned = [batch.shard_indices(n, world, r) for r in range(world)]
            flat = sorted(i for o in owned for i in o)
            assert flat == list(range(n))
            assert max(len(o) for o in owned) - min(len(o) for o in owned) <= 1
    with pytest.raises(ValueError):
        batch.shard_indices(4, 2, 2)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, num_volumes, q):
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        idx = batch.shard_indices(num_volumes, world, rank)
        local = batch.run_shard(idx, lambda i: {"volume": i, "rank": rank, "keypoints": 10 * i + 1})
        out = batch.gather_summaries(local, dist)
        q.put((rank, out))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_batch_gather_world2_gloo():
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    world, n = 2, 7
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=90) for _ in range(world))
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert results[1] is None
    got = results[0]
    assert [g["volume"] for g in got] == list(range(n))
    assert [g["rank"] for g in got] == [i % world for i in range(n)]
    assert [g["keypoints"] for g in got] == [10 * i + 1 for i in range(n)]
