#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path (contract in the task statement).

    python bench.py --gpus N --steps K --warmup W            # our CUDA path
    python bench.py --impl reference --gpus N ...            # the reference's CPU path (oracle port)

Headline (BASELINE.json `metric`, config 2 as written): matched px/s for iter_proj + refine_matches at
512x384, D=24, 10 LM iterations, radius 3, on ONE batch of 64 synthetic frame pairs SHARDED by pair
over the N ranks (strong scaling; pairs are independent, no data-path collective).  A "step" is one
pass of iter_proj + refine_matches over the whole batch with all inputs resident in HBM (3.1 GB over
the N GPUs: >= 390 MB per GPU at N = 8, larger than the 126 MB L2, so every step streams from HBM).
The same JSON line carries:
  * `roofline`   -- the measured matching path (both kernels) on SURVEY.md's 257 algorithmic B/px
                    against the measured HBM copy bandwidth (MEASURED_PEAKS.json); per-kernel numbers
                    under `roofline.kernels`.
  * `e2e`        -- the same metric through the public operator API with inputs in pinned HOST
                    memory (H2D of the four inputs, all kernels, D2H of idx + valid every step).
  * `cpu_baseline` -- the CPU oracle (port of the reference's numpy path, all host threads) timed on
                    a bounded sample of the same batch in the same run (rank 0, N=1 only).
  * `match`      -- parity evidence of THIS run: pairs of the timed B=64 launch checked against the
                    oracle; at N>1 the sharded result compared with the unsharded launch on rank 0.
  * `match_weak` -- 64 pairs PER GPU (the round-1 headline), kept as a second key.
  * `match_full`, `match_single_pair` (config 1), `match_incoherent` (the reference's random fixture).
  * `gn`         -- config 5: Gauss-Newton iterations/s on the 200-keyframe / 1000-edge graph at full
                    196 608 points per keyframe, edges sharded over the N ranks, with bitwise checks
                    against the single-GPU solve, a CPU baseline and an e2e through the L1 API.
  * `gn_rays`, `gn_calib`, `track` -- configs 3 and 4 and the tracker solves (rank 0).
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
PAIRS_TOTAL = 64
MAX_ITER, LAMBDA, THRESH, RADIUS = 10, 1e-8, 1e-6, 3
ALG_BYTES_ITER_PROJ = 36 + 12 + 8 + 8 + 1        # rays + target + p_init in, p + valid out
ALG_BYTES_REFINE = 96 + 96 + 8 + 8               # D11 + D21 + p in, p1 out
ALG_BYTES_PATH = 257                             # fused iter_proj+refine
GN_KF, GN_LOOPS = 200, 103
METRIC = "matched_px_per_s (iter_proj+refine_matches, 512x384, D=24, 10 it, r=3)"
WORKLOAD = ("config 2: one batch of 64 synthetic 512x384 frame pairs, desc dim 24, iter_proj (10 iters) + "
            "refine_matches radius 3, sharded by pair over the GPUs")


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


def bind_to_gpu_numa(local_rank):
    """Pin this rank's host threads to the CPUs NVML reports as local to its GPU, BEFORE any pinned
    buffer is allocated (first touch then places the pages on that NUMA node).  Returns a short
    description for the JSON line."""
    try:
        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(local_rank)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cpus = {64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1}
        cpus &= os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            return {"cpus": len(cpus), "first": min(cpus), "last": max(cpus)}
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)[:80]}
    return {"cpus": 0}


def oracle_threads():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


def cpu_oracle_rate(host, n_pairs_budget_s=12.0):
    """Time the CPU oracle (C port of the reference's numpy path) on a bounded sample of the batch.
    Returns (px_per_s, cores, sample description)."""
    from oracle import oracle
    oracle.build()
    oracle.set_threads(oracle_threads())
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
    ap.add_argument("--pairs", type=int, default=PAIRS_TOTAL, help="pairs of the whole batch per step")
    ap.add_argument("--no-gn", action="store_true", help="skip the Gauss-Newton / tracker legs")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU baseline legs")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extra", action="store_true", help="skip match_weak / single pair / incoherent legs")
    ap.add_argument("--no-check", action="store_true", help="skip the in-run parity checks")
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

    numa = bind_to_gpu_numa(int(os.environ.get("LOCAL_RANK", "0")))
    import numpy as np
    import torch
    from mast3r_slam_b200 import _lib, dist as mdist, kernels, synth

    rank, local_rank, world = mdist.init_from_env()
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    dev = torch.device("cuda", local_rank)
    torch.cuda.set_device(dev)
    _lib.load()
    peak, peak_src = load_peaks()
    BT = args.pairs
    pb, pe = mdist.shard_range(BT, rank, world)
    B = pe - pb                       # this rank's pairs of the one batch
    stream = torch.cuda.current_stream(dev)

    def barrier():
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize()

    # ---- workload: config 2, pair p of the batch has seed p; rank r holds pairs [pb, pe) ----------
    def prepared(seeds):
        ps = synth.make_pairs_torch(seeds, dev, H, W, D, args.desc)
        r, t, p = kernels.prep_for_iter_proj(ps["X11"], ps["X21"], None)
        n = len(list(seeds))
        return dict(rays=r, pts=t, p0=p, D11=ps["D11"], D21=ps["D21"].reshape(n, NPIX, D), X11=ps["X11"],
                    X21=ps["X21"])

    wl = prepared(range(pb, pe))
    torch.cuda.synchronize()

    def timed_match(w, steps, warmup):
        """-> (ms/step max over ranks, iter_proj ms, refine ms, last outputs)"""
        for _ in range(warmup):
            p, valid = kernels.iter_proj(w["rays"], w["pts"], w["p0"], MAX_ITER, LAMBDA, THRESH)
            kernels.refine_matches(w["D11"], w["D21"], p, RADIUS, 0)
        ev = [[torch.cuda.Event(enable_timing=True) for _ in range(3)] for _ in range(steps)]
        barrier()
        for k in range(steps):
            ev[k][0].record(stream)
            p, valid = kernels.iter_proj(w["rays"], w["pts"], w["p0"], MAX_ITER, LAMBDA, THRESH)
            ev[k][1].record(stream)
            p1 = kernels.refine_matches(w["D11"], w["D21"], p, RADIUS, 0)
            ev[k][2].record(stream)
        torch.cuda.synchronize()
        total = mdist.max_over_ranks(ev[0][0].elapsed_time(ev[-1][2]), dev) / steps
        ip = sum(e[0].elapsed_time(e[1]) for e in ev) / steps
        rf = sum(e[1].elapsed_time(e[2]) for e in ev) / steps
        return total, ip, rf, (p, valid, p1)

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t_wall0 = time.perf_counter()
    ms_per_step, t_ip_ms, t_rf_ms, (p_last, valid_last, p1_last) = timed_match(wl, args.steps, args.warmup)
    t_wall = time.perf_counter() - t_wall0
    clocks = sampler.stop() if rank == 0 else None
    value = BT * NPIX / (ms_per_step * 1e-3)

    # ---- parity evidence of this run ---------------------------------------------------------------
    match = {}
    if not args.no_check:
        from oracle import oracle  # the checker, after the timed region
        oracle.build()
        oracle.set_threads(max(1, oracle_threads() // max(world, 1)))
        chk = sorted({0, B - 1}) if B > 0 else []
        ok = True
        for k in chk:
            h = {n: wl[n][k:k + 1].cpu().numpy() for n in ("rays", "pts", "p0", "D11", "D21")}
            po, vo = oracle.iter_proj(h["rays"], h["pts"], h["p0"], MAX_ITER, LAMBDA, THRESH)
            p1o = oracle.refine_matches(h["D11"], h["D21"], po.astype(np.int32), RADIUS, 0)
            ok &= bool(np.array_equal(p_last[k:k + 1].cpu().numpy(), po) and
                       np.array_equal(valid_last[k:k + 1].cpu().numpy().astype(bool), vo) and
                       np.array_equal(p1_last[k:k + 1].cpu().numpy(), p1o))
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=dev)
        if world > 1:
            torch.distributed.all_reduce(flag, op=torch.distributed.ReduceOp.MIN)
        match = {"parity_checked_pairs": [pb + k for k in chk] if world == 1 else
                 f"first and last local pair of every rank ({2 * world} pairs)",
                 "parity_vs_oracle_bit_exact": bool(flag.item()),
                 "launch_checked": f"the timed B={B} launch of each rank (iter_proj p, valid; refine p1)"}
    full0 = None
    if world > 1 and not args.no_check:
        # sharded == unsharded: rank 0 runs the whole 64-pair batch in one launch; every rank's shard
        # of the timed launch is gathered and compared bit for bit
        shard = max(mdist.shard_range(BT, r, world)[1] - mdist.shard_range(BT, r, world)[0] for r in range(world))
        pad_p1 = torch.zeros((shard, NPIX, 2), dtype=torch.int32, device=dev)
        pad_v = torch.zeros((shard, NPIX), dtype=torch.uint8, device=dev)
        pad_p1[:B] = p1_last
        pad_v[:B] = valid_last.to(torch.uint8)
        g_p1 = torch.empty((world * shard, NPIX, 2), dtype=torch.int32, device=dev)
        g_v = torch.empty((world * shard, NPIX), dtype=torch.uint8, device=dev)
        torch.distributed.all_gather_into_tensor(g_p1, pad_p1)
        torch.distributed.all_gather_into_tensor(g_v, pad_v)
        same = torch.ones(1, dtype=torch.int32, device=dev)
        if rank == 0:
            full0 = prepared(range(BT))
            pf, vf = kernels.iter_proj(full0["rays"], full0["pts"], full0["p0"], MAX_ITER, LAMBDA, THRESH)
            p1f = kernels.refine_matches(full0["D11"], full0["D21"], pf, RADIUS, 0)
            for r in range(world):
                b0, e0 = mdist.shard_range(BT, r, world)
                if not (torch.equal(g_p1[r * shard:r * shard + e0 - b0], p1f[b0:e0]) and
                        torch.equal(g_v[r * shard:r * shard + e0 - b0], vf[b0:e0].to(torch.uint8))):
                    same[0] = 0
            del pf, vf, p1f
        torch.distributed.broadcast(same, 0)
        match["sharded_equals_unsharded"] = bool(same.item())
        del g_p1, g_v, pad_p1, pad_v
    del p_last, valid_last, p1_last

    # ---- roofline: the path on 257 B/px, per-kernel numbers below it -------------------------------
    units = B * NPIX
    sm_ghz, sms = 1.965, 148
    kinfo = {
        "iter_proj_tile_kernel": {
            "ms": t_ip_ms, "alg_bytes": ALG_BYTES_ITER_PROJ * units,
            "limiter": "shared-memory crossbar + instruction issue (see DESIGN.md s4, profiles/)",
            "pipe_floor_ms": units * MAX_ITER * 69 / (62.0 * sms * sm_ghz * 1e9) * 1e3,
            "pipe_floor": "69 DADD/DMUL per point-iteration at 62 fp64 lanes/clk/SM (measured)"},
        "refine_tiled_kernel": {
            "ms": t_rf_ms, "alg_bytes": ALG_BYTES_REFINE * units,
            "limiter": "FMA pipe co-limited by the shared-memory crossbar (see DESIGN.md s4, profiles/)",
            "pipe_floor_ms": units * 49 * 24 / (126.0 * sms * sm_ghz * 1e9) * 1e3,
            "pipe_floor": "1176 FMA per pixel at 126 FMA/clk/SM (FFMA2, measured)"},
    }
    for v in kinfo.values():
        v["gbs"] = v["alg_bytes"] / (v["ms"] * 1e-3) / 1e9
    dom = max(kinfo, key=lambda k: kinfo[k]["ms"])
    path_gbs = ALG_BYTES_PATH * units / ((t_ip_ms + t_rf_ms) * 1e-3) / 1e9
    traffic = load_traffic()
    tr = [traffic[k]["dram_bytes_per_px"] for k in kinfo if k in traffic]
    roofline = {
        "bound": "hbm", "kernel": "iter_proj_tile_kernel + refine_tiled_kernel (the measured path; dominant: "
                                  + dom + ")",
        "achieved": round(path_gbs, 1), "peak": peak, "unit": "GB/s", "frac": round(path_gbs / peak, 4),
        "alg_bytes_per_px": ALG_BYTES_PATH,
        "traffic": round(sum(tr) * units) if len(tr) == len(kinfo) else None,
        "peak_source": peak_src, "scope": "this rank's shard of the batch (rank 0)",
        "kernels": {k: {"ms_per_launch": round(v["ms"], 4), "alg_bytes_per_launch": v["alg_bytes"],
                        "achieved_gbs": round(v["gbs"], 1), "frac": round(v["gbs"] / peak, 4),
                        "limiter": v["limiter"], "pipe_floor_ms": round(v["pipe_floor_ms"], 4),
                        "pipe_floor": v["pipe_floor"],
                        "frac_of_pipe_floor": round(v["pipe_floor_ms"] / v["ms"], 4)}
                    for k, v in kinfo.items()},
    }

    # ---- extra: whole match_iterative_proj (prep + iter_proj + refine + tail), ONE C-ABI enqueue ----
    for _ in range(2):
        kernels.match_iterative_proj_device(wl["X11"], wl["X21"], wl["D11"], wl["D21"])
    m0, m1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    m0.record(stream)
    for _ in range(3):
        kernels.match_iterative_proj_device(wl["X11"], wl["X21"], wl["D11"], wl["D21"])
    m1.record(stream)
    torch.cuda.synchronize()
    t_full_ms = mdist.max_over_ranks(m0.elapsed_time(m1) / 3, dev)
    match_full = {"api": "m3s_match_iterative_proj (prep + iter_proj + refine_matches + tail, one enqueue)",
                  "ms_per_step": round(t_full_ms, 4), "value": BT * NPIX / (t_full_ms * 1e-3),
                  "unit": "px/s", "alg_bytes_per_px": 221,
                  "achieved_gbs": round(221 * B * NPIX / (t_full_ms * 1e-3) / 1e9, 1),
                  "frac": round(221 * B * NPIX / (t_full_ms * 1e-3) / 1e9 / peak, 4)}

    # ---- e2e: the reference-facing call on HOST buffers -------------------------------------------
    # matching.match_iterative_proj(X11, X21, D11, D21) (matching.py:339, the call mast3r_utils.py:490
    # makes) with page-locked host inputs and outputs: per step H2D of the four inputs, prep +
    # iter_proj + refine_matches + tail on the device, D2H of idx + valid, chunk-pipelined inside
    # m3s_match_iterative_proj_host.  Synchronous call => host clock between device synchronisations.
    e2e = None
    host = None
    if not args.no_e2e:
        pin = lambda t: torch.empty(t.shape, dtype=t.dtype, pin_memory=True).copy_(t)  # noqa: E731
        hX11, hX21, hD11, hD21 = pin(wl["X11"]), pin(wl["X21"]), pin(wl["D11"]), pin(wl["D21"].reshape(B, H, W, D))
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
        t_mine = 1e3 * (time.perf_counter() - t0) / e_steps
        t_e2e_ms = mdist.max_over_ranks(t_mine, dev)
        rates = torch.tensor([h2d / (t_mine * 1e-3) / 1e9], dtype=torch.float64, device=dev)
        if world > 1:
            allr = torch.empty(world, dtype=torch.float64, device=dev)
            torch.distributed.all_gather_into_tensor(allr, rates)
            rates = allr
        e2e = {"value": BT * NPIX / (t_e2e_ms * 1e-3), "unit": "px/s",
               "h2d_bytes_per_step": h2d * world, "d2h_bytes_per_step": d2h * world,
               "ms_per_step": round(t_e2e_ms, 3), "steps": e_steps,
               "api": "matching.match_iterative_proj on page-locked host arrays (m3s_match_iterative_proj_host)",
               "chunk_pairs": args.chunk_pairs,
               "timer": "host clock around the synchronous call, device synchronised on both sides, max over ranks",
               "h2d_gbs_per_rank": [round(float(v), 1) for v in rates.cpu().tolist()],
               "numa_binding": numa,
               "gpu_launches_per_step": 6 * ((B + args.chunk_pairs - 1) // args.chunk_pairs)}
        if world == 1 and not args.no_cpu:
            host = {"rays": wl["rays"].cpu().numpy(), "pts": wl["pts"].cpu().numpy(), "p0": wl["p0"].cpu().numpy(),
                    "D11": hD11.numpy(), "D21": hD21.numpy().reshape(B, NPIX, D)}
        del hX11, hX21

    # ---- CPU baseline (rank 0, N = 1 only) -------------------------------------------------------
    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu:
        if host is None:
            host = {k: wl[k].cpu().numpy() for k in ("rays", "pts", "p0", "D11", "D21")}
        v, cores, sample = cpu_oracle_rate(host)
        cpu_baseline = {"value": v, "unit": "px/s", "cores": cores, "kind": "port", "sample": sample}
    host = None

    # ---- config 1: one pair, L2 flushed between launches -------------------------------------------
    single = incoherent = weak = None
    if not args.no_extra:
        if rank == 0:
            single = single_pair_leg(wl, dev)
            incoherent = incoherent_leg(dev, peak)
        if world > 1:
            # 64 pairs PER GPU (weak scaling, round 1's headline)
            del wl
            torch.cuda.empty_cache()
            ww = full0 if (rank == 0 and full0 is not None) else prepared(range(1000 * rank, 1000 * rank + BT))
            full0 = None
            w_ms, w_ip, w_rf, _ = timed_match(ww, max(3, min(args.steps, 10)), 3)
            weak = {"pairs_per_gpu": BT, "scaling": "weak", "ms_per_step": round(w_ms, 4),
                    "value": world * BT * NPIX / (w_ms * 1e-3), "unit": "px/s",
                    "iter_proj_ms": round(w_ip, 4), "refine_ms": round(w_rf, 4)}
            del ww
        else:
            weak = {"pairs_per_gpu": BT, "scaling": "weak", "ms_per_step": round(ms_per_step, 4), "value": value,
                    "unit": "px/s", "note": "identical to the headline at N = 1"}
    wl = full0 = None
    torch.cuda.empty_cache()

    # ---- GN legs ---------------------------------------------------------------------------------
    gn = gn_rays = gn_calib = track = None
    if not args.no_gn:
        first = "f64" if args.gn_precision == "f64" else "mixed"
        gn = gn_leg(args, rank, world, dev, peak, first, check=not args.no_check)
        if args.gn_precision == "both":
            f64 = gn_leg(args, rank, world, dev, peak, "f64", check=False)
            gn["pose_diff_vs_f64"] = float((gn.pop("_poses") - f64.pop("_poses")).abs().max())
            gn["gn_f64"] = {k: f64[k] for k in ("value", "ms_per_iter", "accumulate_ms",
                                                "solve_update_ms", "solve_path", "roofline", "status", "precision")}
        gn.pop("_poses", None)
        if rank == 0:
            if world == 1 and not args.no_cpu:
                gn["cpu_baseline"] = gn_cpu_baseline()
            if world == 1 and not args.no_e2e:
                gn["e2e"] = gn_e2e_leg(args, dev)
            gn_rays = gn_rays_leg(dev, peak)
            gn_calib = gn_calib_leg(dev)
            track = track_leg(dev)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "px/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms_per_step, 4), "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 (+f64 bilinear blend)",
            "data": "synthetic",
            "config": {"workload": WORKLOAD, "pairs_total": BT, "pairs_per_gpu": B, "H": H, "W": W, "D": D,
                       "descriptors": args.desc,
                       "l2_policy": f"inputs larger than L2 ({0.0485 * B:.2f} GB per GPU per step vs 126 MB L2)",
                       "parallelism": f"pairs sharded x{world} (no collective)"},
            "clocks": clocks, "e2e": e2e, "gpu_launches": 4 * args.steps, "roofline": roofline,
            "cpu_baseline": cpu_baseline, "match": match, "match_weak": weak, "match_full": match_full,
            "match_single_pair": single, "match_incoherent": incoherent, "gn": gn, "gn_rays": gn_rays,
            "gn_calib": gn_calib, "track": track, "wall_s_timed_region": round(t_wall, 4),
        }
        print(json.dumps(line))
    if world > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


def _flusher(dev):
    """write a buffer larger than L2 (126 MB) between timed launches of a small working set"""
    import torch
    buf = torch.empty(192 * 1024 * 1024, dtype=torch.uint8, device=dev)
    return lambda: buf.add_(1)


def single_pair_leg(wl, dev):
    """Config 1: one 512x384 pair, B = 1 launches (768 CTAs ~ 1.7 waves): device latency per pair with
    the L2 flushed before every launch (the 50 MB working set would otherwise stay L2-resident)."""
    import torch
    from mast3r_slam_b200 import kernels
    one = {k: wl[k][:1].contiguous() for k in ("rays", "pts", "p0", "D11", "D21", "X11", "X21")}
    flush = _flusher(dev)
    st = torch.cuda.current_stream(dev)
    res = {}

    def l1():
        p, _ = kernels.iter_proj(one["rays"], one["pts"], one["p0"], MAX_ITER, LAMBDA, THRESH)
        kernels.refine_matches(one["D11"], one["D21"], p, RADIUS, 0)

    for name, fn in (("iter_proj+refine_matches", l1),
                     ("match_iterative_proj", lambda: kernels.match_iterative_proj_device(
                         one["X11"], one["X21"], one["D11"], one["D21"]))):
        for _ in range(3):
            fn()
        ts = []
        for _ in range(15):
            flush()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(st)
            fn()
            b.record(st)
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        res[name] = {"ms_per_pair_median": round(ts[len(ts) // 2], 4), "ms_min": round(ts[0], 4),
                     "px_per_s": NPIX / (ts[len(ts) // 2] * 1e-3)}
    res["workload"] = "config 1: single synthetic 512x384 pair (seed of pair 0), L2 flushed before each launch"
    return res


def incoherent_leg(dev, peak):
    """SURVEY s8d config 1, last sentence: the reference's own random fixture
    (benchmark_all_kernels.py:67-76 random ray image / targets / uniform p_init, :385-395 unit-norm
    descriptors and uniform centres) at 512x384, full frame, B = 8.  Nothing is spatially coherent, so
    neither tile kernel can stage a useful window: reported separately, with the fraction of queries
    that left the tile paths (m3s_set_path_counters)."""
    import ctypes as C
    import numpy as np
    import torch
    from mast3r_slam_b200 import _lib, kernels, synth
    b = 8
    rays, pts, p0 = synth.reference_like_iter_proj(42, b, H, W, NPIX)
    D11, D21, p1 = synth.reference_like_refine(43, b, H, W, D, NPIX)
    t = lambda a: torch.from_numpy(a).to(dev)  # noqa: E731
    rays, pts, p0, D11, D21, p1 = t(rays), t(pts), t(p0), t(D11), t(D21), t(p1)
    lib = _lib.load()
    cnt = torch.zeros(3, dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream(dev)
    for _ in range(2):
        p, _ = kernels.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
        kernels.refine_matches(D11, D21, p1, RADIUS, 0)
    _lib.check(lib.m3s_set_path_counters(C.c_void_p(cnt.data_ptr())))
    try:
        kernels.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
        kernels.refine_matches(D11, D21, p1, RADIUS, 0)
        torch.cuda.synchronize()
    finally:
        _lib.check(lib.m3s_set_path_counters(None))
    c = cnt.cpu().tolist()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    n = 5
    ev[0].record(st)
    for _ in range(n):
        kernels.iter_proj(rays, pts, p0, MAX_ITER, LAMBDA, THRESH)
    ev[1].record(st)
    for _ in range(n):
        kernels.refine_matches(D11, D21, p1, RADIUS, 0)
    ev[2].record(st)
    torch.cuda.synchronize()
    ip, rf = ev[0].elapsed_time(ev[1]) / n, ev[1].elapsed_time(ev[2]) / n
    units = b * NPIX
    gbs = ALG_BYTES_PATH * units / ((ip + rf) * 1e-3) / 1e9
    return {"workload": "reference's random fixture (benchmark_all_kernels.py:67-76, 385-395) at 512x384, "
                        "full frame, 8 pairs per launch; refine centres are the fixture's uniform p1",
            "value": units / ((ip + rf) * 1e-3), "unit": "px/s", "iter_proj_ms": round(ip, 4),
            "refine_ms": round(rf, 4), "achieved_gbs": round(gbs, 1), "frac": round(gbs / peak, 4),
            "refine_queries_off_tile_frac": round(c[0] / units, 4),
            "iter_proj_points_without_window_frac": round(c[1] / units, 4),
            "iter_proj_points_left_window_frac": round(c[2] / units, 4)}


def gn_leg(args, rank, world, dev, peak, precision, check=True):
    """Config 5: gauss_newton_points on a 200-KF / 1000-directed-edge graph, N = 196 608."""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib, dist as mdist, kernels, synth
    e_total = 2 * ((GN_KF - 1) + (GN_KF - 2) + GN_LOOPS)
    eb, ee = mdist.shard_edge_range(e_total, rank, world, 2)
    do_check = check and world > 1
    # rank 0 materialises every edge row when it also has to solve the whole graph on one GPU
    g = synth.make_graph_torch(dev, 0, GN_KF, GN_LOOPS, H, W,
                               edge_range=None if (do_check and rank == 0) else (eb, ee))
    assert len(g["ii"]) == e_total
    params = kernels.make_params(_lib.GN_POINTS, args.gn_iters, 1, 0.0, 0.0, 1.5, 0.01,
                                 precision=precision)
    Twc = torch.from_numpy(g["Twc"]).to(dev)
    lib = _lib.load()
    stream = torch.cuda.current_stream(dev)
    st = C.c_void_p(stream.cuda_stream)
    flags = {}
    full = None
    if do_check and rank == 0:
        full = g
        single = kernels.GNPlan(g["ii"], g["jj"], GN_KF, NPIX, 1, dev)
        _lib.check(lib.m3s_gn_set_poses(single.handle, C.c_void_p(Twc.data_ptr()), st))
        pp = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
        for _ in range(args.gn_iters):
            _lib.check(lib.m3s_gn_accumulate(single.handle, C.byref(params), pp(g["Xs"]), pp(g["Cs"]), pp(g["idx"]),
                                             pp(g["valid"]), pp(g["Q"]), 0, e_total, st))
            _lib.check(lib.m3s_gn_solve_update(single.handle, C.byref(params), st))
        torch.cuda.synchronize()
        single_poses = single.poses().clone()
        del single
        g = dict(g)
        g["idx"], g["valid"], g["Q"] = (full[k][eb:ee].contiguous() for k in ("idx", "valid", "Q"))
        full = None
        torch.cuda.empty_cache()
    sg = mdist.ShardedGaussNewton(g["ii"], g["jj"], GN_KF, NPIX, 1, dev, eb, ee)

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
    poses_after = sg.plan.poses().clone()
    if do_check:
        # every rank compares its fp64 pose state after the timed iterations with rank 0's
        # single-GPU solve of the whole graph, bit for bit
        ref = poses_after.clone() if rank != 0 else single_poses
        torch.distributed.broadcast(ref, 0)
        eq = torch.tensor([1 if torch.equal(ref, poses_after) else 0], dtype=torch.int32, device=dev)
        torch.distributed.all_reduce(eq, op=torch.distributed.ReduceOp.MIN)
        mine = poses_after.clone()
        torch.distributed.broadcast(mine, 0)    # rank 0's SHARDED result
        ag = torch.tensor([1 if torch.equal(mine, poses_after) else 0], dtype=torch.int32, device=dev)
        torch.distributed.all_reduce(ag, op=torch.distributed.ReduceOp.MIN)
        flags = {"bitwise_vs_single_gpu": bool(eq.item()), "ranks_agree": bool(ag.item()),
                 "checked": f"fp64 pose state [200,8] after the {args.gn_iters} timed iterations, every rank"}
    # accumulate kernel alone (dominant): event-timed on its stream
    a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reset()
    sync()
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
    err0 = float(abs(g["Twc"] - g["Twc_gt"]).max())
    exchange = sg.exchange
    solve_path = int(lib.m3s_gn_plan_factor_path(sg.plan.handle))
    sparse_info = sg.plan.sparse_info()
    sg.close()
    out = {
        "metric": "gn_iters_per_s (gauss_newton_points, 200 KF / 1000 directed edges, 196608 pts/KF)",
        "value": 1e3 / ms, "unit": "iter/s", "ms_per_iter": round(ms, 4), "iters_timed": args.gn_iters,
        "n_gpus": world, "scaling": "strong (edges sharded; packed H,g blocks exchanged once per iteration)",
        "exchange": {"p2p": "peer stores from gn_edge_finalize_kernel over NVLink (CUDA IPC) + epoch flags",
                     "nccl": "NCCL all-reduce of the packed blocks", "none": "single GPU"}[exchange],
        "edges_this_rank": ne, "dim": sg.plan.dim,
        "accumulate_ms": round(acc_ms, 4), "solve_update_ms": round(ms - acc_ms, 4),
        "solve_path": {0: "steps (one launch per block step)", 1: "fused (persistent chain + workers)",
                       2: "tiles (dataflow over shared-memory-resident tiles)",
                       3: "sparse front (levels of exactly eliminated keyframes) + tiles on the dense tail"}[solve_path],
        "sparse_front": sparse_info,
        "precision": precision, "_poses": poses_after.cpu(),
        "roofline": {"bound": "hbm",
                     "kernel": "gn_accum_stream_kernel" if precision == "mixed" else "gn_accum_kernel",
                     "alg_bytes_per_launch": alg,
                     "achieved": round(alg / (acc_ms * 1e-3) / 1e9, 1), "peak": peak, "unit": "GB/s",
                     "frac": round(alg / (acc_ms * 1e-3) / 1e9 / peak, 4),
                     "iteration_frac": round(alg / (ms * 1e-3) / 1e9 / peak, 4)},
        "status": status,
        "pose_err_before": err0,
    }
    out.update(flags)
    return out


def gn_cpu_baseline():
    """The oracle port (C/OpenMP restatement of gauss_newton_points.py, all host threads) on the SAME
    200-KF / 1000-edge topology at num_pts' = 2048 and 4096 sub-sampled points per keyframe; the
    per-iteration time is linear in the points (gauss_newton_points.py:99-160) plus a fixed solve, so
    the two runs give slope and intercept and the full-size figure is their extrapolation."""
    from mast3r_slam_b200 import synth
    from oracle import oracle
    oracle.build()
    oracle.set_threads(oracle_threads())
    ts = {}
    for n in (2048, 4096):
        g = synth.make_graph(0, GN_KF, GN_LOOPS, H, W, num_pts=n)
        a = (g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
        oracle.gauss_newton_points(*a, max_iter=1, delta_thresh=0.0)
        t0 = time.perf_counter()
        oracle.gauss_newton_points(*a, max_iter=2, delta_thresh=0.0)
        ts[n] = (time.perf_counter() - t0) / 2
    slope = (ts[4096] - ts[2048]) / 2048.0
    fixed = max(ts[2048] - slope * 2048, 0.0)
    full = fixed + slope * NPIX
    return {"value": 1.0 / full, "unit": "iter/s", "cores": oracle.num_threads(), "kind": "port",
            "sample": f"1000 edges x 2048 and x 4096 points per keyframe, 2 iterations each "
                      f"({ts[2048]:.3f} s, {ts[4096]:.3f} s per iteration); extrapolated linearly in points to "
                      f"196608 (+ {fixed:.3f} s fixed solve): {full:.1f} s per iteration"}


def gn_e2e_leg(args, dev):
    """kernels.gauss_newton_points (the L1 operator, kernels.py:396-455) on HOST numpy arrays: the
    call uploads Twc / Xs / Cs / idx / valid / Q (2.4 GB), runs the solver loop, returns numpy poses."""
    import numpy as np
    import torch
    from mast3r_slam_b200 import kernels, synth
    g = synth.make_graph_torch(dev, 0, GN_KF, GN_LOOPS, H, W)
    host = {k: g[k].cpu().numpy() for k in ("Xs", "Cs", "idx", "Q")}
    host["valid"] = g["valid"].cpu().numpy().astype(bool)
    ii, jj, Twc = g["ii"], g["jj"], g["Twc"]
    del g
    torch.cuda.empty_cache()
    h2d = sum(v.nbytes for v in host.values()) + Twc.nbytes
    it = args.gn_iters

    def call():
        return kernels.gauss_newton_points(Twc, host["Xs"], host["Cs"], ii, jj, host["idx"], host["valid"],
                                           host["Q"], max_iter=it, delta_thresh=0.0)

    call()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = call()
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    assert isinstance(out, np.ndarray)
    return {"value": it / t, "unit": "iter/s", "iters_per_call": it, "s_per_call": round(t, 4),
            "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": int(out.nbytes),
            "api": "kernels.gauss_newton_points on pageable host numpy arrays (one call = upload + "
                   f"{it} iterations + download)"}


def _time_gn_calls(fn, reps, dev):
    import torch
    for _ in range(2):
        fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    a.record()
    for _ in range(reps):
        out = fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps, out


def gn_rays_leg(dev, peak):
    """Config 3: gauss_newton_rays on a 10-keyframe / 40-directed-edge graph at full 196 608 points,
    device-resident inputs, through kernels.gauss_newton_rays (10 iterations, delta 1e-4, pin 1)."""
    import torch
    from mast3r_slam_b200 import kernels, synth
    g = synth.make_graph_torch(dev, 3, 10, 3, H, W)
    Twc = torch.from_numpy(g["Twc"]).to(dev)
    a = (Twc, g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
    it = 5
    ms, out = _time_gn_calls(lambda: kernels.gauss_newton_rays(*a, max_iter=it, delta_thresh=0.0), 5, dev)
    e = len(g["ii"])
    alg = 9 * e * NPIX + 16 * 10 * NPIX + (7 * 9) ** 2 * 8
    err0 = float(abs(g["Twc"] - g["Twc_gt"]).max())
    err1 = float((out.cpu() - torch.from_numpy(g["Twc_gt"])).abs().max())
    return {"workload": f"config 3: gauss_newton_rays, 10 KF / {e} directed edges, 196608 pts/KF, {it} iterations per call",
            "ms_per_iter": round(ms / it, 4), "value": 1e3 * it / ms, "unit": "iter/s",
            "alg_bytes_per_iter": alg, "achieved_gbs": round(alg / (ms / it * 1e-3) / 1e9, 1),
            "frac": round(alg / (ms / it * 1e-3) / 1e9 / peak, 4),
            "l2_note": "102 MB per iteration fits the 126 MB L2: iterations after the first re-read from L2",
            "pose_err_before": err0, "pose_err_after": err1}


def gn_calib_leg(dev):
    """Config 4: tracking-style gauss_newton_calib solves on a synthetic TUM-like sequence: per frame a
    2-keyframe problem {keyframe pinned, frame free}, E = 2 (both directions), 196 608 points, known
    intrinsics, sigma_pixel 1, sigma_depth 10."""
    import torch
    from mast3r_slam_b200 import kernels, synth
    frames = 8
    gs = [synth.make_graph_torch(dev, 100 + f, 2, 0, H, W) for f in range(frames)]
    K4 = gs[0]["K4"]
    prob = []
    for g in gs:
        # (constrain_points_to_ray is not applied: the synthetic keyframes store their points in a
        # circularly shifted order, so point n does not sit on pixel n's ray)
        prob.append((torch.from_numpy(g["Twc"]).to(dev), g["Xs"], g["Cs"], K4, g["ii"], g["jj"], g["idx"],
                     g["valid"], g["Q"]))

    def solve_all():
        outs = []
        for a in prob:
            outs.append(kernels.gauss_newton_calib(*a, (W, H), 0, 0.0, 1.0, 10.0, max_iter=10, delta_thresh=1e-4))
        return outs

    ms, outs = _time_gn_calls(solve_all, 3, dev)
    err0 = max(float(abs(g["Twc"] - g["Twc_gt"]).max()) for g in gs)
    err1 = max(float((o.cpu() - torch.from_numpy(g["Twc_gt"])).abs().max()) for o, g in zip(outs, gs))
    return {"workload": f"config 4: gauss_newton_calib, {frames} frames x (2 KF, 2 directed edges, 196608 pts), "
                        "max_iter 10, delta 1e-4, device-resident inputs",
            "ms_per_frame_solve": round(ms / frames, 4), "value": 1e3 * frames / ms, "unit": "solves/s",
            "pose_err_before": err0, "pose_err_after": err1}


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
    vectors (tests/golden).  Each step = the whole 64-pair batch of our arm (`--pairs`).  Rank 0 alone
    runs; no GPU is touched."""
    if rank != 0:
        return
    import numpy as np
    from mast3r_slam_b200 import synth
    from oracle import oracle
    oracle.build()
    # torchrun exports OMP_NUM_THREADS=1 for every rank; the reference arm is a CPU measurement and
    # uses every host core this process may run on
    oracle.set_threads(oracle_threads())
    k = args.pairs
    ps = synth.make_pairs(range(k), H, W, D, args.desc)
    rays, pts, p0 = oracle.prep_for_iter_proj(ps["X11"], ps["X21"], None)
    D11, D21 = ps["D11"], ps["D21"].reshape(k, NPIX, D)
    del ps
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
    sample = f"all {k} pairs per step, iter_proj+refine_matches, {t / args.steps:.3f} s/step"
    print(json.dumps({
        "impl": "reference", "metric": METRIC,
        "value": v, "unit": "px/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": round(1e3 * t / args.steps, 3), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32 (+f64 bilinear blend)", "data": "synthetic",
        "config": {"workload": WORKLOAD, "pairs_total": k, "pairs_per_gpu": k, "H": H, "W": W, "D": D,
                   "descriptors": args.desc},
        "cpu_baseline": {"value": v, "unit": "px/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "px/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }))


if __name__ == "__main__":
    main()
