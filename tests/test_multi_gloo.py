"""CPU tests of the N > 1 host logic (no GPU, no compute):
  * bench.py's multi-process plumbing -- one process per GPU, barrier + max-over-ranks timing,
    whole-job value = units of all ranks / slowest rank's time -- on a world-size-2 gloo group;
  * the z-slab layout rule of sift3d_detect_describe_sharded (host C, sift3d_sharded_layout)."""
import ctypes as C
import os
import socket

import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        # every rank "times" its own K steps; the job's time is the slowest rank's
        ms = 3.0 + rank  # rank 1 is slower
        t = torch.tensor([ms], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        host = dist.new_group(backend="gloo")  # the host-side barrier bench.py uses around C5
        dist.barrier(group=host)
        q.put((rank, float(t[0])))
    finally:
        dist.destroy_process_group()


def test_max_over_ranks_timing_and_host_barrier_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
    assert got == {0: 4.0, 1: 4.0}
    assert all(p.exitcode == 0 for p in procs)
    # whole-job throughput as bench.py forms it: units of all ranks / max-over-ranks time
    nvox, steps, world = 256 ** 3, 10, 2
    value = world * nvox * steps / (got[0] * steps * 1e-3)
    assert value == pytest.approx(2 * nvox / 4.0e-3)


def test_sharded_layout_rule():
    """owned ranges tile the volume on multiples of 2^K; every cut face carries a halo that
    covers the widest window of the path at the coarsest sharded octave"""
    import sift3d_b200 as s3
    L = s3.lib()
    for nz, num, K in ((1024, 8, 4), (1024, 8, 3), (250, 2, 1), (480, 3, 2), (512, 8, 3)):
        arr = [(C.c_int * num)() for _ in range(4)]
        halo = L.sift3d_sharded_layout(nz, num, K, 40, 8, 1.0, *arr)
        g0, g1, a0, a1 = [list(a) for a in arr]
        assert halo > 0 and halo % (1 << K) == 0 and halo >= 40 << (K - 1)
        assert g0[0] == 0 and g1[-1] == nz and all(g1[r] == g0[r + 1] for r in range(num - 1))
        assert all(g % (1 << K) == 0 for g in g0) and all(g1[r] > g0[r] for r in range(num))
        assert all(a0[r] == max(0, g0[r] - halo) and a1[r] == min(nz, g1[r] + halo) for r in range(num))
    # anisotropic: the filter's reach in planes grows as the spacing shrinks
    arr = [(C.c_int * 2)() for _ in range(4)]
    assert L.sift3d_sharded_layout(512, 2, 1, 10, 8, 0.25, *arr) >= 34
    arr = [(C.c_int * 4)() for _ in range(4)]
    assert L.sift3d_sharded_layout(24, 4, 3, 40, 8, 1.0, *arr) == -1
