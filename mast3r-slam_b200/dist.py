"""Multi-GPU partitioning of the hot path: one process per GPU, torch.distributed for plumbing.

Two shardings, nothing else:
  * matching  -- frame pairs are independent: the batch dimension is split contiguously over ranks,
                 NO collective on the data path.
  * Gauss-Newton -- directed-edge rows are split contiguously over ranks; Xs / Cs / poses are
                 replicated.  Each iteration every rank reduces its own edges into packed per-edge
                 blocks ([E, 36] fp64) and the ranks exchange them:
                   exchange="p2p"  (default on CUDA) -- gn_edge_finalize_kernel stores each owned
                     block directly into every rank's receive buffer over NVLink peer mappings
                     (CUDA IPC handles exchanged once through torch.distributed) and publishes an
                     epoch flag; a wait kernel acquires the flags.  No collective call per iteration.
                   exchange="nccl" -- slots a rank does not own are zero, ONE all-reduce(sum) fills
                     them in (x + 0 is exact).
                 Either way the buffer is bit-identical to the single-GPU one; every rank then
                 assembles H, g, solves and retracts redundantly -- identical inputs, identical
                 poses, no broadcast.
                 The buffer is the block-sparse form of the 7F x 7F Hessian and gradient:
                 E*36*8 B (288 KB for 1000 edges) instead of 15.5 MB dense.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np
import torch
import torch.distributed as dist

from . import _lib, kernels


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous, balanced [begin, end) of `total` units for `rank` (first total % world ranks get
    one extra).  Covers [0, total) exactly once over all ranks."""
    base, rem = divmod(total, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def shard_edge_range(num_edges: int, rank: int, world: int, group_size: int = 1) -> tuple[int, int]:
    """Edge shard that keeps `group_size` consecutive rows together (2 when both directions of a
    pair are adjacent rows)."""
    groups = (num_edges + group_size - 1) // group_size
    b, e = shard_range(groups, rank, world)
    return min(b * group_size, num_edges), min(e * group_size, num_edges)


def init_from_env(backend: str | None = None):
    """torchrun-style rendezvous (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*).  Returns
    (rank, local_rank, world).  Single-process when WORLD_SIZE is unset or 1."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            dist.init_process_group(backend, device_id=torch.device("cuda", local_rank))
        else:
            dist.init_process_group(backend)
    return rank, local_rank, world


def allreduce_edge_blocks(blocks: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the packed per-edge blocks over ranks in place (NCCL on CUDA tensors, gloo on CPU)."""
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(blocks, op=dist.ReduceOp.SUM, group=group)
    return blocks


def max_over_ranks(value: float, device=None) -> float:
    """Max of a per-rank scalar (timings are reported as the slowest rank's)."""
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64,
                     device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def match_sharded(X11, X21, D11, D21, rank: int, world: int, **kw):
    """Pairs split by rank; returns this rank's (begin, end), idx, valid.  No collective."""
    b, e = shard_range(X11.shape[0], rank, world)
    idx, valid = kernels.match_iterative_proj_device(X11[b:e], X21[b:e], D11[b:e], D21[b:e], **kw)
    return (b, e), idx, valid


class ShardedGaussNewton:
    """Edge-sharded Gauss-Newton over a process group.

    `idx`, `valid`, `Q` hold only this rank's rows [edge_begin, edge_end) of the directed-edge
    arrays (shape [edge_end - edge_begin, N]); `ii`, `jj` describe the WHOLE graph on every rank.
    """

    def __init__(self, ii, jj, num_kf, num_pts, pin, device, edge_begin, edge_end, group=None,
                 exchange: str | None = None):
        # a private plan (not the per-graph cache of kernels.get_plan): attaching the peer exchange
        # changes what m3s_gn_accumulate does, which single-GPU callers of a shared plan must not see
        ii_np = np.ascontiguousarray(np.asarray(ii, dtype=np.int32).reshape(-1))
        jj_np = np.ascontiguousarray(np.asarray(jj, dtype=np.int32).reshape(-1))
        self.plan = kernels.GNPlan(ii_np, jj_np, num_kf, num_pts, pin, device)
        self.device = device
        self.edge_begin, self.edge_end = int(edge_begin), int(edge_end)
        self.group = group
        self.blocks = self.plan.edge_blocks()
        self._lib = _lib.load()
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        if exchange is None:
            exchange = os.environ.get("M3S_GN_EXCHANGE", "p2p")
        if exchange not in ("p2p", "nccl"):
            raise ValueError(f"exchange must be 'p2p' or 'nccl', got {exchange!r}")
        self.exchange = exchange if world > 1 else "none"
        if self.exchange == "p2p":
            self._attach_peers(world)

    def _attach_peers(self, world: int):
        """Exchange the CUDA IPC handles of the per-rank receive regions (host plumbing, once per
        plan) and map the peers.  Raises on every rank if any rank cannot map its peers."""
        lib, plan = self._lib, self.plan
        rank = dist.get_rank(self.group)
        torch.cuda.synchronize(self.device)   # detach + re-create must not race in-flight work
        handle = (C.c_ubyte * 64)()
        _lib.check(lib.m3s_gn_exchange_create(plan.handle, rank, world, handle), "gn_exchange_create")
        mine = torch.tensor(list(bytes(handle)), dtype=torch.uint8, device=self.device)
        every = torch.empty(world * 64, dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(every, mine, group=self.group)
        buf = (C.c_ubyte * (world * 64)).from_buffer_copy(bytes(every.cpu().numpy().tobytes()))
        rc = lib.m3s_gn_exchange_attach(plan.handle, buf)
        ok = torch.tensor([1 if rc == 0 else 0], dtype=torch.int32, device=self.device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)   # doubles as the barrier
        if int(ok.item()) != 1:
            msg = lib.m3s_last_error().decode() if rc != 0 else "a peer rank failed to attach"
            lib.m3s_gn_exchange_detach(plan.handle)
            raise _lib.M3SError(f"gn peer exchange unavailable ({msg}); pass exchange='nccl'")

    def close(self):
        """Unmap the peers (collective: every rank calls it)."""
        if self.exchange == "p2p":
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)
            self._lib.m3s_gn_exchange_detach(self.plan.handle)
            dist.barrier(group=self.group)   # every peer has unmapped before any region is freed
            self._lib.m3s_gn_exchange_release(self.plan.handle)
            self.exchange = "closed"

    def _local_base(self, t, row_elems, itemsize):
        """device pointer such that row `edge_begin` of the virtual full array is t[0]"""
        return C.c_void_p(t.data_ptr() - self.edge_begin * row_elems * itemsize)

    def iterate(self, params, Xs, Cs, idx, valid, Q, iters: int):
        """Enqueue `iters` Gauss-Newton iterations (no host synchronisation)."""
        lib, plan = self._lib, self.plan
        n = plan.N
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        for _ in range(iters):
            _lib.check(lib.m3s_gn_accumulate(
                plan.handle, C.byref(params), p(Xs), p(Cs), self._local_base(idx, n, 4),
                self._local_base(valid, n, 1), self._local_base(Q, n, 4), self.edge_begin,
                self.edge_end, st), "gn_accumulate")
            if self.exchange == "nccl":
                allreduce_edge_blocks(self.blocks, self.group)
            elif self.exchange == "closed":
                raise _lib.M3SError("ShardedGaussNewton used after close()")
            _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), st), "gn_solve_update")

    def solve(self, params, Twc, Xs, Cs, idx, valid, Q):
        """Full solver: poses [K,8] f32 CUDA in -> poses out (all ranks return the same bits)."""
        lib, plan = self._lib, self.plan
        st = C.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        _lib.check(lib.m3s_gn_set_poses(plan.handle, C.c_void_p(Twc.data_ptr()), st))
        if plan.F > 0:
            self.iterate(params, Xs, Cs, idx, valid, Q, params.max_iter)
        out = torch.empty_like(Twc)
        _lib.check(lib.m3s_gn_get_poses(plan.handle, C.c_void_p(out.data_ptr()), st))
        return out

    def check(self):
        """Synchronise and raise if the last solve failed; an exchange timeout (a peer that did not
        publish within M3S_GN_XCH_TIMEOUT_MS) is reported as such, not as a numerical failure."""
        st = self.plan.status()
        if st["exchange_timeout"]:
            raise _lib.M3SError("gn peer exchange timed out waiting for a peer rank "
                                "(M3S_GN_XCH_TIMEOUT_MS; 0 disables the bound)")
        if st["failed"]:
            raise _lib.M3SError("gauss-newton solve failed (non-finite or not positive definite)")
        return st
