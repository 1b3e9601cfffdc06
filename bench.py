#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (contract in the task statement).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --gpus N ...            # the reference's CPU path (oracle port)

Headline (BASELINE.json `metric`): matched px/s for iter_proj + refine_matches at 512x384, D=24,
10 LM iterations, radius 3, on a batch of 64 synthetic frame pairs per GPU (config 2; weak scaling:
pairs are independent, no data-path collective).  A "step" is one pass of iter_proj + refine_matches
over the whole 64-pair batch with all inputs resident in HBM (3.1 GB per step: far larger than the
126 MB L2, so every step streams from HBM).  The same JSON line carries:
  * `e2e`       -- the same metric through the public operator API with inputs in pinned HOST
                   memory: per step H2D of all five inputs, both kernels, D2H of the refined
                   positions and validity.
  * `roofline`  -- for the dominant kernel: algorithmic bytes per launch / CUDA-event duration
                   against the measured HBM copy bandwidth (MEASURED_PEAKS.json).
  * `cpu_baseline` -- the CPU oracle (port of the reference's numpy path, all host threads) timed
                   on a bounded sample of the same batch in the same run (rank 0, N=1 only).
  * `gn`        -- the second half of BASELINE.json's metric: Gauss-Newton iterations/s on the
                   200-keyframe / 1000-edge graph (config 5) at full 196 608 points per keyframe,
                   gauss_newton_points, edges sharded over the N ranks with one NCCL all-reduce of
                   the packed (H, g) blocks per iteration (strong scaling of one fixed graph).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

H, W, D = 384, 512, 24
NPIX = H * W
PAIRS_PER_GPU = 64
MAX_ITER, LAMBDA, THRESH, RADIUS = 10, 1e-8, 1e-6, 3
ALG_BYTES_ITER_PROJ = 36 + 12 + 8 + 8 + 1        # rays + target + p_init in, p + valid out
ALG_BYTES_REFINE = 96 + 96 + 8 + 8               # D11 + D21 + p in, p1 out
ALG_BYTES_PATH = 257                             # fused iter_proj+refine
GN_KF, GN_LOOPS = 200, 103


def load_traffic():
    """measured DRAM traffic per unit (ncu --set full captures, profiles/traffic.json)"""
    try:
        return json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
    except Exception:
        return {}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows, self.proc, self.gpu = [], None, gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms",
                 "200", "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            pass
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for k, nm in enumerate(names):
                    if r[5 + k].lower().startswith("active"):
                        reasons.add(nm)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def cpu_oracle_rate(host, n_pairs_budget_s=12.0, threads=None):
    """Time the CPU oracle (C port of the reference's numpy path) on a bounded sample of the batch.
    Returns (px_per_s, cores, sample description)."""
    from oracle import oracle
    oracle.build()
    if threads:
        oracle.set_threads(threads)
    cores = oracle.num_threads()

    def run(k):
        t0 = time.perf_counter()
        p, _ = oracle.iter_proj(host["rays"][:k], host["pts"][:k], host["p0"][:k], MAX_ITER, LAMBDA, THRESH)
        oracle.refine_matches(host["D11"][:k], host["D21"][:k], p.astype("int32"), RADIUS, 0)
        return time.perf_counter() - t0

    t1 = run(1)
    k = int(max(1, min(host["rays"].shape[0], n_pairs_budget_s / max(t1, 1e-3))))
    t = run(k) if k > 1 else t1
    return k * NPIX / t, cores, f"{k} of the {host['rays'].shape[0]} pairs (first k), iter_proj+refine_matches, {t:.2f} s"


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pairs", type=int, default=PAIRS_PER_GPU, help="pairs per GPU per step")
    ap.add_argument("--no-gn", action="store_true", help="skip the 200-keyframe Gauss-Newton leg")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline leg")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--chunk-pairs", type=int, default=4, help="pairs per pipelined chunk (e2e leg)")
    ap.add_argument("--gn-iters", type=int, default=5)
    ap.add_argument("--gn-precision", default="both", choices=["both", "mixed", "f64"],
                    help="per-point arithmetic of the GN accumulate pass (include/m3s.h); 'both' "
                         "reports the default (mixed) as `gn` and the fp64 pass as `gn_f64`")
    ap.add_argument("--desc", default="dyadic")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        return reference_arm(args, rank, world)

    import numpy as np
    import torch
    from mast3r_slam_b200 import _lib, dist as mdist, kernels, synth

    rank, local_rank, world = mdist.init_from_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    lib = _lib.load()
    peak, peak_src = load_peaks()
    B = args.pairs

    # ---- workload: config 2, seeds rank*B .. rank*B+B-1 (weak scaling: B pairs per GPU) ----------
    pairs = synth.make_pairs_torch(range(rank * B, rank * B + B), dev, H, W, D, args.desc)
    rays, pts, p0 = kernels.prep_for_iter_proj(pairs["X11"], pairs["X21"], None)
    D11 = pairs["D11"]
    D21 = pairs["D21"].reshape(B, NPIX, D)
    X11, X21 = pairs["X11"], pairs["X21"]
    del pairs
    torch.cuda.synchronize()

    def step_device():
        p, valid = kernels.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
        p1 = kernels.refine_matches(D11, D21, p, RADIUS, 0)
        return p1, valid

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(args.steps)]
    stream = torch.cuda.current_stream(dev)
    barrier()
    t_wall0 = time.perf_counter()
    for k in range(args.steps):
        ev[k][0].record(stream)
        p, valid = kernels.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
        ev[k][1].record(stream)
        p1 = kernels.refine_matches(D11, D21, p, RADIUS, 0)
        ev[k][2].record(stream)
    torch.cuda.synchronize()
    t_wall = time.perf_counter() - t_wall0
    t_total_ms = ev[0][0].elapsed_time(ev[-1][2])
    t_ip_ms = sum(e[0].elapsed_time(e[1]) for e in ev) / args.steps
    t_rf_ms = sum(e[1].elapsed_time(e[2]) for e in ev) / args.steps
    clocks = sampler.stop() if rank == 0 else None
    t_max_ms = mdist.max_over_ranks(t_total_ms, dev)
    ms_per_step = t_max_ms / args.steps
    value = world * B * NPIX / (ms_per_step * 1e-3)

    # ---- extra: the whole match_iterative_proj (prep + iter_proj + refine + tail) as
    # ONE C-ABI enqueue from X11, X21, D11, D21 resident in HBM; 221 algorithmic B/px ------------
    for _ in range(2):
        kernels.match_iterative_proj_device(X11, X21, D11, D21)
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    m0.record(stream)
    for _ in range(3):
        kernels.match_iterative_proj_device(X11, X21, D11, D21)
    m1.record(stream)
    torch.cuda.synchronize()
    t_full_ms = mdist.max_over_ranks(m0.elapsed_time(m1) / 3, dev)
    match_full = {"api": "m3s_match_iterative_proj (prep + iter_proj + refine_matches + tail, one enqueue)",
                  "ms_per_step": round(t_full_ms, 4), "value": world * B * NPIX / (t_full_ms * 1e-3),
                  "unit": "px/s", "alg_bytes_per_px": 221,
                  "achieved_gbs": round(221 * B * NPIX / (t_full_ms * 1e-3) / 1e9, 1),
                  "frac": round(221 * B * NPIX / (t_full_ms * 1e-3) / 1e9 / peak, 4)}

    # ---- roofline of the dominant kernel (per launch = one 64-pair batch) ------------------------
    units = B * NPIX
    kinfo = {
        "iter_proj_tile_kernel": {"ms": t_ip_ms, "alg_bytes": ALG_BYTES_ITER_PROJ * units},
        "refine_tiled_kernel": {"ms": t_rf_ms, "alg_bytes": ALG_BYTES_REFINE * units},
    }
    # what actually bounds each kernel (ncu, profiles/): neither is HBM-bound, so the HBM fraction is
    # reported next to the SM-pipe floor of the exact arithmetic
    sm_ghz, sms = 1.965, 148
    kinfo["iter_proj_tile_kernel"].update(
        limiter="shared-memory crossbar: 288 B of fp64 corners per point-iteration (ncu s17: LSU wavefronts "
                "86 % of peak, issue 66 %, fp64 pipe 31 %, XU 34 %, DRAM 6 %)",
        pipe_floor_ms=units * MAX_ITER * 69 / (62.0 * sms * sm_ghz * 1e9) * 1e3,
        pipe_floor="69 DADD/DMUL per point-iteration at 62 fp64 lanes/clk/SM (measured)")
    kinfo["refine_tiled_kernel"].update(
        limiter="FMA issue at 8 warps/SM (ncu s17: FFMA2 = 56 % of the hot loop, FMA pipe 40 %, LSU 58 %, DRAM 19 %)",
        pipe_floor_ms=units * 49 * 24 / (126.0 * sms * sm_ghz * 1e9) * 1e3,
        pipe_floor="1176 FMA per pixel at 126 FMA/clk/SM (FFMA2, measured)")
    for v in kinfo.values():
        v["gbs"] = v["alg_bytes"] / (v["ms"] * 1e-3) / 1e9
    dom = max(kinfo, key=lambda k: kinfo[k]["ms"])
    path_gbs = ALG_BYTES_PATH * units / ((t_ip_ms + t_rf_ms) * 1e-3) / 1e9
    roofline = {
        "bound": "hbm", "kernel": dom, "achieved": round(kinfo[dom]["gbs"], 1), "peak": peak,
        "unit": "GB/s", "frac": round(kinfo[dom]["gbs"] / peak, 4),
        "traffic": (round(load_traffic()[dom]["dram_bytes_per_px"] * units) if dom in load_traffic() else None),
        "peak_source": peak_src,
        "kernels": {k: {"ms_per_launch": round(v["ms"], 4), "alg_bytes_per_launch": v["alg_bytes"],
                        "achieved_gbs": round(v["gbs"], 1), "frac": round(v["gbs"] / peak, 4),
                        "limiter": v["limiter"], "pipe_floor_ms": round(v["pipe_floor_ms"], 4),
                        "pipe_floor": v["pipe_floor"],
                        "frac_of_pipe_floor": round(v["pipe_floor_ms"] / v["ms"], 4)}
                    for k, v in kinfo.items()},
        "path": {"alg_bytes_per_px": ALG_BYTES_PATH, "achieved_gbs": round(path_gbs, 1),
                 "frac": round(path_gbs / peak, 4)},
    }

    # ---- e2e: the reference-facing call on HOST buffers -------------------------------------------
    # matching.match_iterative_proj(X11, X21, D11, D21) (matching.py:339, the call mast3r_utils.py:490
    # makes) with page-locked host inputs and outputs: per step H2D of the four inputs, prep +
    # iter_proj + refine_matches + tail on the device, D2H of idx + valid, chunk-pipelined inside
    # m3s_match_iterative_proj_host.  The call is synchronous, so it is timed with the host clock
    # between device synchronisations.  `l1_unpipelined` = the same through the two L1 operator
    # calls (rays / pts / p_init / D11 / D21 uploaded whole, one stream), for comparison.
    e2e = None
    host = None
    if not args.no_e2e:
        pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)  # noqa: E731
        hX11, hX21, hD11, hD21 = pin(X11), pin(X21), pin(D11), pin(D21.reshape(B, H, W, D))
        out_idx = torch.empty((B, NPIX), dtype=torch.int32, pin_memory=True)
        out_val = torch.empty((B, NPIX), dtype=torch.uint8, pin_memory=True)
        h2d = sum(t.numel() * t.element_size() for t in (hX11, hX21, hD11, hD21))
        d2h = out_idx.numel() * 4 + out_val.numel()

        def step_e2e():
            kernels.match_iterative_proj_host(hX11, hX21, hD11, hD21, None, MAX_ITER, LAMBDA, THRESH,
                                              RADIUS, 0, 0.1, chunk_pairs=args.chunk_pairs,
                                              out=(out_idx, out_val), device=dev)

        e_steps = max(2, min(args.steps, 5))
        step_e2e()
        barrier()
        t0 = time.perf_counter()
        for _ in range(e_steps):
            step_e2e()
        torch.cuda.synchronize()
        t_e2e_ms = mdist.max_over_ranks(1e3 * (time.perf_counter() - t0), dev) / e_steps
        e2e = {"value": world * B * NPIX / (t_e2e_ms * 1e-3), "unit": "px/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "ms_per_step": round(t_e2e_ms, 3),
               "steps": e_steps, "api": "matching.match_iterative_proj on page-locked host arrays "
               "(m3s_match_iterative_proj_host)", "chunk_pairs": args.chunk_pairs,
               "timer": "host clock around the synchronous call, device synchronised on both sides",
               "gpu_launches_per_step": 6 * ((B + args.chunk_pairs - 1) // args.chunk_pairs)}
        # L1 operators, whole-batch upload on one stream
        hp = {k: pin(t) for k, t in (("rays", rays), ("pts", pts), ("p0", p0))}
        out_p1 = torch.empty((B, NPIX, 2), dtype=torch.int32, pin_memory=True)

        def step_l1():
            d = {k: t.to(dev, non_blocking=True) for k, t in hp.items()}
            d11, d21 = hD11.to(dev, non_blocking=True), hD21.to(dev, non_blocking=True)
            p, valid = kernels.iter_proj(d["rays"], d["pts"], d["p0"], MAX_ITER, LAMBDA, THRESH)
            p1 = kernels.refine_matches(d11, d21.reshape(B, NPIX, D), p, RADIUS, 0)
            out_p1.copy_(p1, non_blocking=True)
            out_val.copy_(valid, non_blocking=True)

        step_l1()
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(2):
            step_l1()
        e1.record(stream)
        torch.cuda.synchronize()
        t_l1 = mdist.max_over_ranks(e0.elapsed_time(e1), dev) / 2
        e2e["l1_unpipelined"] = {
            "value": world * B * NPIX / (t_l1 * 1e-3), "ms_per_step": round(t_l1, 3),
            "h2d_bytes_per_step": h2d - hX11.numel() * 8 + sum(t.numel() * 4 for t in hp.values()),
            "d2h_bytes_per_step": out_p1.numel() * 4 + out_val.numel()}
        host = {"rays": hp["rays"].numpy(), "pts": hp["pts"].numpy(), "p0": hp["p0"].numpy(),
                "D11": hD11.numpy(), "D21": hD21.numpy().reshape(B, NPIX, D)}
        del hX11, hX21

    # ---- CPU baseline (rank 0, N = 1 only) -------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        if host is None:
            host = {k: t.cpu().numpy() for k, t in (("rays", rays), ("pts", pts), ("p0", p0), ("D11", D11), ("D21", D21))}
        v, cores, sample = cpu_oracle_rate(host)
        cpu_baseline = {"value": v, "unit": "px/s", "cores": cores, "kind": "port", "sample": sample}
    del rays, pts, p0, D11, D21, X11, X21
    host = None
    torch.cuda.empty_cache()

    # ---- GN leg: 200-KF graph, edges sharded over ranks ------------------------------------------
    gn = None
    if not args.no_gn:
        first = "f64" if args.gn_precision == "f64" else "mixed"
        gn = gn_leg(args, rank, world, dev, peak, first)
        if args.gn_precision == "both":
            f64 = gn_leg(args, rank, world, dev, peak, "f64")
            gn["pose_diff_vs_f64"] = float((gn.pop("_poses") - f64.pop("_poses")).abs().max())
            gn["gn_f64"] = {k: f64[k] for k in ("value", "ms_per_iter", "accumulate_ms",
                                                "solve_update_ms", "roofline", "status", "precision")}
        gn.pop("_poses", None)

    track = None
    if rank == 0 and not args.no_gn:
        track = track_leg(dev)

    if rank == 0:
        line = {
            "metric": "matched_px_per_s (iter_proj+refine_matches, 512x384, D=24, 10 it, r=3)",
            "value": value, "unit": "px/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 (+f64 bilinear blend)",
            "data": "synthetic",
            "config": {"workload": f"config 2: batch of {B} synthetic 512x384 frame pairs per GPU, "
                                   "desc dim 24, iter_proj (10 iters) + refine_matches radius 3",
                       "pairs_per_gpu": B, "H": H, "W": W, "D": D, "descriptors": args.desc,
                       "l2_policy": "inputs larger than L2 (3.1 GB per step vs 126 MB L2)",
                       "parallelism": f"pairs x{world} (no collective)"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 4 * args.steps, "roofline": roofline,
            "cpu_baseline": cpu_baseline, "match_full": match_full, "gn": gn, "track": track, "wall_s_timed_region": round(t_wall, 4),
        }
        print(json.dumps(line))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def gn_leg(args, rank, world, dev, peak, precision):
    """Config 5: gauss_newton_points on a 200-KF / 1000-directed-edge graph, N = 196 608."""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib, dist as mdist, kernels, synth
    e_total = 2 * ((GN_KF - 1) + (GN_KF - 2) + GN_LOOPS)
    eb, ee = mdist.shard_edge_range(e_total, rank, world, 2)
    g = synth.make_graph_torch(dev, 0, GN_KF, GN_LOOPS, H, W, edge_range=(eb, ee))
    assert len(g["ii"]) == e_total
    sg = mdist.ShardedGaussNewton(g["ii"], g["jj"], GN_KF, NPIX, 1, dev, eb, ee)
    params = kernels.make_params(_lib.GN_POINTS, args.gn_iters, 1, 0.0, 0.0, 1.5, 0.01,
                                 precision=precision)
    Twc = torch.from_numpy(g["Twc"]).to(dev)
    lib = _lib.load()
    stream = torch.cuda.current_stream(dev)
    st = C.c_void_p(stream.cuda_stream)

    def reset():
        _lib.check(lib.m3s_gn_set_poses(sg.plan.handle, C.c_void_p(Twc.data_ptr()), st))

    def sync():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    reset()
    sg.iterate(params, g["Xs"], g["Cs"], g["idx"], g["valid"], g["Q"], 2)  # warm-up
    sync()
    reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    e0.record(stream)
    sg.iterate(params, g["Xs"], g["Cs"], g["idx"], g["valid"], g["Q"], args.gn_iters)
    e1.record(stream)
    torch.cuda.synchronize()
    ms = mdist.max_over_ranks(e0.elapsed_time(e1), dev) / args.gn_iters
    status = sg.plan.status()
    poses_after = torch.empty_like(Twc)
    _lib.check(lib.m3s_gn_get_poses(sg.plan.handle, C.c_void_p(poses_after.data_ptr()), st))
    torch.cuda.synchronize()
    # accumulate kernel alone (dominant): event-timed on its stream
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reset()
    torch.cuda.synchronize()
    a0.record(stream)
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    for _ in range(3):
        _lib.check(lib.m3s_gn_accumulate(
            sg.plan.handle, C.byref(params), p(g["Xs"]), p(g["Cs"]), sg._local_base(g["idx"], NPIX, 4),
            sg._local_base(g["valid"], NPIX, 1), sg._local_base(g["Q"], NPIX, 4), eb, ee, st))
    a1.record(stream)
    torch.cuda.synchronize()
    acc_ms = a0.elapsed_time(a1) / 3
    ne = ee - eb
    alg = 9 * ne * NPIX + 16 * GN_KF * NPIX + (7 * 199) ** 2 * 8
    out = torch.empty_like(Twc)
    _lib.check(lib.m3s_gn_get_poses(sg.plan.handle, p(out), st))
    torch.cuda.synchronize()
    err0 = float(abs(g["Twc"] - g["Twc_gt"]).max())
    exchange = sg.exchange
    sg.close()
    return {
        "metric": "gn_iters_per_s (gauss_newton_points, 200 KF / 1000 directed edges, 196608 pts/KF)",
        "value": 1e3 / ms, "unit": "iter/s", "ms_per_iter": round(ms, 4), "iters_timed": args.gn_iters,
        "n_gpus": world, "scaling": "strong (edges sharded; packed H,g blocks exchanged once per iteration)",
        "exchange": {"p2p": "peer stores from gn_edge_finalize_kernel over NVLink (CUDA IPC) + epoch flags",
                     "nccl": "NCCL all-reduce of the packed blocks", "none": "single GPU"}[exchange],
        "edges_this_rank": ne, "dim": sg.plan.dim,
        "accumulate_ms": round(acc_ms, 4), "solve_update_ms": round(ms - acc_ms, 4),
        "precision": precision, "_poses": poses_after.double().cpu(),
        "roofline": {"bound": "hbm",
                     "kernel": "gn_accum_stream_kernel" if precision == "mixed" else "gn_accum_kernel",
                     "alg_bytes_per_launch": alg,
                     "achieved": round(alg / (acc_ms * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(alg / (acc_ms * 1e-3) / 1e9 / peak, 4)},
        "status": status,
        "pose_err_before": err0,
    }


def track_leg(dev):
    """one frame-to-keyframe Sim3 solve (tracker.py:255-406) at 512x384, inputs resident
    on the device, event-timed over 20 solves (each = init + up to 10 x (accumulate + solve))."""
    import torch
    from mast3r_slam_b200 import synth, tracker
    d = synth.make_track_pair(0, H, W)
    t = lambda a, dt=torch.float32: torch.from_numpy(a).to(dev, dt)  # noqa: E731
    Qk, v_opt, _, (n_opt, _) = tracker.track_gate(t(d["Cf"]), t(d["Ck"]), t(d["Qff"]), t(d["Qkf"]),
                                                  t(d["idx_f2k"], torch.int32),
                                                  t(d["valid_match"].astype("uint8"), torch.uint8))
    Xf, Xk, idx = t(d["Xf"]), t(d["Xk"]), t(d["idx_f2k"], torch.int32)
    Tf, Tk = t(d["T_WCf"]), t(d["T_WCk"])
    out = {}
    for name, fn in (("ray_dist", lambda: tracker.opt_pose_ray_dist_sim3(Xf, Xk, Tf, Tk, Qk, v_opt, idx_f2k=idx)),
                     ("calib", lambda: tracker.opt_pose_calib_sim3(Xf, Xk, Tf, Tk, Qk, v_opt, d["K4"], (H, W),
                                                                    idx_f2k=idx))):
        for _ in range(3):
            fn()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(20):
            res = fn()
        e1.record()
        torch.cuda.synchronize()
        out[name] = {"ms_per_solve_api": round(e0.elapsed_time(e1) / 20, 4),
                     "pose_err_after": float((res[0].cpu() - torch.from_numpy(d["T_WCf_gt"])).abs().max())}
    # the same solve enqueued back to back through the C ABI on preallocated buffers (no Python
    # tensor plumbing between launches): the device time of one cooperative launch
    import ctypes as C
    from mast3r_slam_b200 import _lib
    lib = _lib.load()
    vq, vv = Qk.reshape(-1).contiguous(), v_opt.reshape(-1).to(torch.uint8).contiguous()
    ws = torch.empty(lib.m3s_track_workspace_bytes(NPIX), dtype=torch.uint8, device=dev)
    o1, o2 = torch.empty(8, device=dev), torch.empty(8, device=dev)
    stats = torch.empty(4, dtype=torch.float64, device=dev)
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for name, mode in (("ray_dist", _lib.TRACK_RAY_DIST), ("calib", _lib.TRACK_CALIB)):
        prm = tracker._params(mode, tracker.get_config()["tracking"], d["K4"], (H, W))
        call = lambda: _lib.check(lib.m3s_track_opt_pose(  # noqa: E731
            C.byref(prm), p(Xf), p(idx), p(Xk), p(vq), p(vv), NPIX, p(Tf), p(Tk), p(o1), p(o2), p(stats),
            p(ws), ws.numel(), st))
        for _ in range(3):
            call()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(50):
            call()
        e1.record()
        torch.cuda.synchronize()
        out[name]["ms_per_solve"] = round(e0.elapsed_time(e1) / 50, 4)
        out[name]["iterations"] = int(stats.cpu()[0])
    out["points"] = NPIX
    out["valid_points"] = n_opt
    out["pose_err_before"] = float(abs(d["T_WCf"] - d["T_WCf_gt"]).max())
    return out


def reference_arm(args, rank, world):
    """The reference's own CPU implementation of the path (its numpy fallbacks, which is what its
    dispatcher executes without Metal), as the C/OpenMP port in oracle/ on all host threads.  The
    reference is pure Python and does not travel to the GPU box; oracle/ is pinned to it by golden
    vectors (tests/golden).  Each step = a bounded sample (`--pairs`, default 2) of the 64-pair
    workload.  Rank 0 alone runs."""
    if rank != 0:
        return
    import numpy as np
    from mast3r_slam_b200 import synth
    from oracle import oracle
    oracle.build()
    # torchrun exports OMP_NUM_THREADS=1 for every rank; the reference arm is a CPU measurement and
    # uses every host core this process may run on
    try:
        oracle.set_threads(len(os.sched_getaffinity(0)))
    except Exception:
        oracle.set_threads(os.cpu_count() or 1)
    k = args.pairs if args.pairs != PAIRS_PER_GPU else 8
    ps = synth.make_pairs(range(k), H, W, D, args.desc)
    rays, pts, p0 = oracle.prep_for_iter_proj(ps["X11"], ps["X21"], None)
    D11, D21 = ps["D11"], ps["D21"].reshape(k, NPIX, D)
    cores = oracle.num_threads()

    def step():
        p, _ = oracle.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
        oracle.refine_matches(D11, D21, p.astype(np.int32), RADIUS, 0)

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    t = time.perf_counter() - t0
    v = k * NPIX * args.steps / t
    sample = f"{k} of the 64 pairs per step, iter_proj+refine_matches, {t / args.steps:.3f} s/step"
    print(json.dumps({
        "impl": "reference",
        "metric": "matched_px_per_s (iter_proj+refine_matches, 512x384, D=24, 10 it, r=3)",
        "value": v, "unit": "px/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(1e3 * t / args.steps, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32 (+f64 bilinear blend)", "data": "synthetic",
        "config": {"workload": "config 2: batch of 64 synthetic 512x384 frame pairs per GPU, desc dim "
                               "24, iter_proj (10 iters) + refine_matches radius 3",
                   "sample_pairs_per_step": k, "H": H, "W": W, "D": D, "descriptors": args.desc},
        "cpu_baseline": {"value": v, "unit": "px/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "px/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


if __name__ == "__main__":
    main()
