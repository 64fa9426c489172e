"""z-slab sharding of one volume on the GPU: the slab path must give
the single-device result bit for bit. The ranks of these tests share one GPU (gloo process
group, halos staged through the host), so they run on the 1-GPU box; tools/slab_check.py is the
same check over NCCL with one GPU per rank."""
import os
import socket
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, dims, units, octaves, params, q):
    import torch.distributed as dist

    import slab_check

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rep = slab_check.run_check(dims, units, octaves, 0, params=params)
        q.put((rank, rep))
    except Exception as e:  # noqa: BLE001 - reported to the parent instead of a queue timeout
        q.put((rank, {"error": repr(e)}))
        raise
    finally:
        dist.destroy_process_group()


def _run(world, dims, units, octaves, params=None):
    import torch.multiprocessing as mp

    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, dims, units, octaves, params, q))
             for r in range(world)]
    for p in procs:
        p.start()
    results = dict(q.get(timeout=280) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
    errors = {r: v["error"] for r, v in results.items() if isinstance(v, dict) and "error" in v}
    assert not errors, errors
    assert all(p.exitcode == 0 for p in procs)
    return results[0]


@pytest.mark.timeout(300)
def test_seeded_tail_octaves_match_whole_pyramid():
    """world 1: no cut faces; octaves >= K run through sift3d_cuda_detect_from_seed."""
    rep = _run(1, (96, 96, 128), (1.0, 1.0, 1.0), 1)
    assert rep["tail_octaves"] >= 1 and rep["candidates"][0] > 50
    assert rep["identical"], rep


@pytest.mark.timeout(300)
@pytest.mark.parametrize("world,dims,octaves", [
    (2, (96, 96, 320), 2),     # two sharded octaves + replicated tail
    (2, (128, 64, 250), 1),    # nz not a multiple of the alignment, one sharded octave
    (3, (64, 64, 480), 2),     # an interior slab with two cut faces
])
def test_slab_path_is_bit_identical(world, dims, octaves):
    rep = _run(world, dims, (1.0, 1.0, 1.0), octaves)
    assert rep["candidates"][0] > 50 and rep["keypoints"][0] > 10
    assert rep["identical"], rep


@pytest.mark.timeout(300)
def test_slab_path_with_parameters_and_coarse_units():
    rep = _run(2, (64, 64, 384), (2.0, 2.0, 2.0), 2,
               {"peak_thresh": 0.05, "corner_thresh": 0.3, "num_kp_levels": 4})
    assert rep["identical"], rep


@pytest.mark.timeout(300)
def test_slab_path_with_anisotropic_units():
    """Non-dyadic tap spacing: the z coordinates z - d/unit are formed in volume coordinates."""
    rep = _run(2, (64, 64, 320), (0.9, 0.9, 1.4), 1)
    assert rep["candidates"][0] > 20
    assert rep["identical"], rep
