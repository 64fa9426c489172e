"""Host logic of the z-slab path (one tall volume over several GPUs)
on CPU: layout, halo exchange and seed gather over world_size-2/3 gloo, candidate merge and
in-order compaction. No GPU and no oracle involved."""
import os
import socket

import numpy as np
import pytest

from sift3d_b200 import slab
from sift3d_b200.api import CUDA_KEYPOINT_DTYPE


def test_required_halo_covers_descriptor_window():
    # defaults: sd of the top keypoint level = 1.6 * 2^(2/3); window 2 * 7.07 * sd = 35.9 voxels
    h = slab.required_halo(1.6, 3, 1.0, 1, 8)
    assert h >= 38 and h % 2 == 0
    h2 = slab.required_halo(1.6, 3, 1.0, 2, 8)
    assert h2 % 4 == 0 and h2 >= 2 * 38
    # finer voxels widen the window in planes
    assert slab.required_halo(1.6, 3, 0.5, 1, 8) >= 2 * 36
    # wide filters dominate when the descriptor window is small
    assert slab.required_halo(0.2, 3, 1.0, 1, 60) >= 60


@pytest.mark.parametrize("nz,world,K", [(256, 2, 2), (250, 3, 1), (1024, 8, 3), (97, 2, 2), (64, 1, 2)])
def test_layout_partitions_every_octave(nz, world, K):
    halo = 8 << K
    layout = slab.slab_layout(nz, world, K, halo)
    assert len(layout) == world
    for o in range(K + 1):
        planes = nz >> o
        covered = []
        for s in layout:
            p0, p1 = s.owned(o)
            covered += list(range(p0, p1))
            h0, h1 = s.held(o)
            assert h0 <= p0 < p1 <= h1 <= planes
            # slab-array extent halves like a volume of a1 - a0 planes (imutil.c:1545-1548)
            assert h1 - h0 == (s.a1 - s.a0) >> o
            # cut faces keep the whole halo
            if s.g0 > 0:
                assert p0 - h0 == min(halo, s.g0) >> o
            own0, own1 = s.own(o)
            assert (own0, own1) == (p0 - h0, p1 - h0)
        assert covered == list(range(planes))
    for s in layout[1:]:
        assert s.g0 % (1 << K) == 0


def test_layout_rejects_impossible_splits():
    with pytest.raises(ValueError):
        slab.slab_layout(16, 8, 2, 8)
    with pytest.raises(ValueError):
        slab.slab_layout(256, 2, 2, 6)


@pytest.mark.parametrize("nz,world,K,halo", [(256, 2, 2, 80), (300, 4, 1, 40), (128, 4, 2, 80)])
def test_exchange_plan_fills_exactly_the_halo(nz, world, K, halo):
    layout = slab.slab_layout(nz, world, K, halo)
    for o in range(K):
        rng = np.random.default_rng(o)
        full = rng.standard_normal((nz >> o, 3, 4)).astype(np.float32)
        held = []
        for s in layout:
            h0, h1 = s.held(o)
            a = np.full((h1 - h0, 3, 4), np.nan, dtype=np.float32)
            p0, p1 = s.owned(o)
            a[p0 - h0:p1 - h0] = full[p0:p1]
            held.append(a)
        for src, dst, p0, p1 in slab.exchange_plan(layout, o):
            bs, bd = layout[src].held(o)[0], layout[dst].held(o)[0]
            seg = held[src][p0 - bs:p1 - bs]
            assert not np.isnan(seg).any()           # only owned planes are ever sent
            assert np.isnan(held[dst][p0 - bd:p1 - bd]).all()  # each plane is written once
            held[dst][p0 - bd:p1 - bd] = seg
        for s, a in zip(layout, held):
            h0, h1 = s.held(o)
            np.testing.assert_array_equal(a, full[h0:h1])


def _synthetic_candidates(rng, dims, noct, kp):
    """A candidate list in the reference's scan order (o, s, z, y, x)."""
    rows = []
    for o in range(noct):
        nx, ny, nz = (d >> o for d in dims)
        for s in range(kp):
            n = int(rng.integers(5, 40))
            z = np.sort(rng.integers(1, nz - 1, n))
            for zz in z:
                rows.append((o, s, float(rng.integers(1, nx - 1)), float(rng.integers(1, ny - 1)),
                             float(zz), 1.6 * 2 ** (o + s / kp), float(rng.random()),
                             rng.standard_normal(9).astype(np.float32), int(rng.random() < 0.4)))
    out = np.zeros(len(rows), dtype=CUDA_KEYPOINT_DTYPE)
    for i, r in enumerate(rows):
        out[i] = r
    return out


def test_merge_and_compaction_reproduce_single_device_order():
    rng = np.random.default_rng(7)
    dims, K, total = (64, 64, 256), 2, 4
    layout = slab.slab_layout(dims[2], 3, K, 16)
    ref = _synthetic_candidates(rng, dims, total, 3)
    per_rank = []
    for s in layout:
        lo = np.array([s.owned(int(o))[0] for o in ref["o"]])
        hi = np.array([s.owned(int(o))[1] for o in ref["o"]])
        per_rank.append(ref[(ref["o"] < K) & (ref["z"] >= lo) & (ref["z"] < hi)])
    tail = ref[ref["o"] >= K]
    merged, order = slab.merge_candidates(per_rank, tail)
    assert len(merged) == len(ref)
    for f in ref.dtype.names:  # (padding bytes of the aligned record are not preserved)
        np.testing.assert_array_equal(merged[f], ref[f])
    assert sorted(order.tolist()) == list(range(len(ref)))

    kps, keep = slab.compact_keypoints(merged)
    assert (merged["valid"][keep] == 1).all() and len(keep) == int(ref["valid"].sum())
    # survivor j carries the strength of candidate j (sift.c:372-384, 1148-1166)
    np.testing.assert_array_equal(kps["strength"], ref["strength"][:len(keep)])
    np.testing.assert_array_equal(kps["z"], ref["z"][keep])


def test_merge_handles_empty_ranks():
    empty = np.zeros(0, dtype=CUDA_KEYPOINT_DTYPE)
    merged, order = slab.merge_candidates([empty, empty], None)
    assert len(merged) == 0 and len(order) == 0
    kps, keep = slab.compact_keypoints(merged)
    assert len(kps) == 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, nz, K, halo, q):
    import torch
    import torch.distributed as dist

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        layout = slab.slab_layout(nz, world, K, halo)
        ok = True
        for o in range(K + 1):
            g = torch.Generator().manual_seed(100 + o)
            full = torch.randn((nz >> o, 5, 6), generator=g)
            h0, h1 = layout[rank].held(o)
            p0, p1 = layout[rank].owned(o)
            mine = torch.full((h1 - h0, 5, 6), float("nan"))
            mine[p0 - h0:p1 - h0] = full[p0:p1]
            if o < K:
                slab.exchange_halos(mine, layout, o, rank)
                ok = ok and torch.equal(mine, full[h0:h1])
            else:  # the seed of the replicated octaves
                whole = torch.full_like(full, float("nan"))
                slab.gather_owned(mine, layout, o, rank, whole)
                ok = ok and torch.equal(whole, full)
        t = torch.tensor([float(rank), 5.0 - rank])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = ok and t.tolist() == [float(world - 1), 5.0]
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.timeout(180)
@pytest.mark.parametrize("world,nz,K,halo", [(2, 96, 2, 16), (3, 200, 1, 80)])
def test_halo_exchange_and_seed_gather_gloo(world, nz, K, halo):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, nz, K, halo, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=30)
        assert p.exitcode == 0
    assert results == {r: True for r in range(world)}
