"""Multi-GPU check (run under torchrun on >= 2 GPUs; not collected by pytest):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 \
        --master-port 29511 tests/mgpu_check.py

1. edge-sharded Gauss-Newton (packed (H, g) blocks exchanged once per iteration, by peer stores
   from the finalize kernel over NVLink or by one NCCL all-reduce) must
   return exactly the poses of the single-GPU solve, on every rank (bitwise), and match the oracle;
2. pair-sharded matching must equal the un-sharded result on the rank's slice.
"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_b200 import _lib, dist as mdist, kernels, synth  # noqa: E402


def main():
    rank, local_rank, world = mdist.init_from_env()
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    g = synth.make_graph(11, 24, 8, 96, 128)
    e_total = len(g["ii"])
    n = g["Xs"].shape[1]
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    single = kernels.gauss_newton_points(Twc, Xs, Cs, g["ii"], g["jj"], idx, valid, Q, max_iter=6)
    eb, ee = mdist.shard_edge_range(e_total, rank, world, 2)
    params = kernels.make_params(_lib.GN_POINTS, 6, 1, 1e-4, 0.0, 1.5, 0.01)
    timing = {}
    for exchange in ("nccl", "p2p", "p2p"):   # p2p twice: a second plan re-creates the mappings
        sg = mdist.ShardedGaussNewton(g["ii"], g["jj"], 24, n, 1, dev, eb, ee, exchange=exchange)
        assert sg.exchange == exchange
        args = (params, Twc, Xs, Cs, idx[eb:ee].contiguous(), valid[eb:ee].contiguous(),
                Q[eb:ee].contiguous())
        sharded = sg.solve(*args)
        torch.cuda.synchronize()
        assert torch.equal(single, sharded), f"rank {rank}: {exchange}-sharded GN differs from single-GPU GN"
        st = sg.plan.status()
        assert not st["failed"], st
        # repeated solves on the same plan (epochs keep counting) + device time per solve
        torch.distributed.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            again = sg.solve(*args)
        e1.record()
        torch.cuda.synchronize()
        assert torch.equal(single, again)
        timing[exchange] = mdist.max_over_ranks(e0.elapsed_time(e1) / 10, dev)
        sg.close()
    # early stop (the delta test fires after a few iterations, the remaining enqueued iterations are
    # no-ops that still run the exchange protocol): consecutive solves on ONE plan, alternating
    # start poses, must stay bit-identical to the single-GPU solver -- the epoch / parity
    # double-buffer must survive solves that stop at different iterations
    Twc2 = Twc.clone()
    Twc2[1:, :3] += 0.003
    stop_params = kernels.make_params(_lib.GN_POINTS, 8, 1, 2e-3, 0.0, 1.5, 0.01)
    ref_stop = [kernels.gauss_newton_points(T, Xs, Cs, g["ii"], g["jj"], idx, valid, Q, max_iter=8,
                                            delta_thresh=2e-3) for T in (Twc, Twc2)]
    st1 = kernels.get_plan(g["ii"], g["jj"], 24, n, 1, dev).status()
    assert st1["stopped"] and 0 < st1["applied"] < 8, f"test graph must stop early: {st1}"
    for exchange in ("p2p", "nccl"):
        sg = mdist.ShardedGaussNewton(g["ii"], g["jj"], 24, n, 1, dev, eb, ee, exchange=exchange)
        sl = (idx[eb:ee].contiguous(), valid[eb:ee].contiguous(), Q[eb:ee].contiguous())
        for rep in range(6):
            k = rep % 2
            out = sg.solve(stop_params, (Twc, Twc2)[k], Xs, Cs, *sl)
            if rep % 3 == 2:
                torch.cuda.synchronize()   # some solves back to back, some separated
            assert torch.equal(out, ref_stop[k]), f"rank {rank}: {exchange} early-stop solve {rep} differs"
        st2 = sg.check()
        assert st2["stopped"] and st2["applied"] == st1["applied"], (st1, st2)
        sg.close()
    # a rank that owns no edges still takes part in the exchange (publishes from the wait kernel)
    eb0, ee0 = (0, e_total) if rank == 0 else (e_total, e_total)
    sg = mdist.ShardedGaussNewton(g["ii"], g["jj"], 24, n, 1, dev, eb0, ee0, exchange="p2p")
    lop = sg.solve(params, Twc, Xs, Cs, idx[eb0:ee0].contiguous(), valid[eb0:ee0].contiguous(),
                   Q[eb0:ee0].contiguous())
    torch.cuda.synchronize()
    assert torch.equal(single, lop), f"rank {rank}: lopsided p2p sharding differs"
    sg.close()
    gathered = [torch.empty_like(sharded) for _ in range(world)]
    torch.distributed.all_gather(gathered, sharded)
    assert all(torch.equal(gathered[0], x) for x in gathered), "ranks disagree"
    if rank == 0:
        from oracle import oracle
        ref = oracle.gauss_newton_points(g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"],
                                         g["Q"], max_iter=6)
        upd = np.abs(ref - g["Twc"]).max()
        err = np.abs(sharded.cpu().numpy().astype(np.float64) - ref).max()
        assert err <= 1e-5 * upd + 2e-7, (err, upd)
    ps = synth.make_pairs(range(4), 48, 64)
    (b, e), idx_s, valid_s = mdist.match_sharded(ps["X11"], ps["X21"], ps["D11"], ps["D21"], rank, world)
    idx_f, valid_f = kernels.match_iterative_proj_device(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    assert np.array_equal(idx_s, idx_f[b:e]) and np.array_equal(valid_s, valid_f[b:e])
    torch.distributed.barrier()
    if rank == 0:
        print(f"mgpu_check ok: world={world}, sharded GN (nccl and p2p exchange) bitwise == single GPU "
              f"(incl. 6 consecutive early-stopped solves per exchange, stop after {st1['applied']} of 8), "
              f"pairs sharded {world}x; ms per 6-iteration solve: " +
              ", ".join(f"{k} {v:.3f}" for k, v in timing.items()))
    torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
