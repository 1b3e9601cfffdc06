"""Host-side logic of the multi-GPU path on CPU: world_size-2 gloo process group.

The CUDA kernels cannot run here; what is covered is everything around them: contiguous pair /
edge sharding, the all-reduce of the packed per-edge (H, g) blocks (each rank fills only the slots
of its own edges, x + 0 is exact, so the reduced buffer equals the single-rank one bit for bit), and
the max-over-ranks timing reduction bench.py uses."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, E, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from mast3r_slam_b200 import dist as mdist
    r, lr, w = mdist.init_from_env("gloo")
    assert (r, w) == (rank, world)
    # the "single GPU" per-edge blocks every rank would compute for its own edges
    full = torch.from_numpy(np.random.default_rng(0).standard_normal((E, 36)))
    b, e = mdist.shard_edge_range(E, rank, world, 2)
    mine = torch.zeros_like(full)
    mine[b:e] = full[b:e]
    mdist.allreduce_edge_blocks(mine)
    ok_blocks = torch.equal(mine, full)  # bitwise
    t = mdist.max_over_ranks(10.0 + rank)
    pb, pe = mdist.shard_range(64, rank, world)
    ret[rank] = (ok_blocks, t, (b, e), (pb, pe))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("E", [1000, 14])
def test_gloo_world2_edge_allreduce(E):
    world = 2
    port = _free_port()
    ctx = mp.get_context("spawn")
    with ctx.Manager() as mgr:
        ret = mgr.dict()
        procs = [ctx.Process(target=_worker, args=(r, world, port, E, ret)) for r in range(world)]
        for p in procs:
            p.start()
        for p in procs:
            p.join(120)
            assert p.exitcode == 0
        res = dict(ret)
    assert all(res[r][0] for r in range(world)), "all-reduced edge blocks differ from single-rank"
    assert all(res[r][1] == 10.0 + world - 1 for r in range(world))
    edges = [res[r][2] for r in range(world)]
    assert edges[0][0] == 0 and edges[-1][1] == E and edges[0][1] == edges[1][0]
    assert edges[0][1] % 2 == 0  # both directions of a pair stay on one rank
    assert [res[r][3] for r in range(world)] == [(0, 32), (32, 64)]


def test_shard_ranges_cover_exactly():
    from mast3r_slam_b200.dist import shard_edge_range, shard_range
    for total in (0, 1, 7, 64, 1000, 1001):
        for world in (1, 2, 3, 4, 8):
            got = [shard_range(total, r, world) for r in range(world)]
            assert got[0][0] == 0 and got[-1][1] == total
            assert all(got[i][1] == got[i + 1][0] for i in range(world - 1))
            sizes = [e - b for b, e in got]
            assert max(sizes) - min(sizes) <= 1
            ge = [shard_edge_range(total, r, world, 2) for r in range(world)]
            assert ge[0][0] == 0 and ge[-1][1] == total
            assert all(ge[i][1] == ge[i + 1][0] for i in range(world - 1))
