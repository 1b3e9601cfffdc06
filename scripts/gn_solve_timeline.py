#!/usr/bin/env python
"""Timeline of the 200-keyframe damped solve (dim 1393, 44 blocks of 32): per-kernel %globaltimer
stamps of the two factorisation chains (diag / panel+syrk) and the dataflow back-substitution.
    M3S_GN_TIMELINE=1 python scripts/gn_solve_timeline.py > gpurun_out/gn_timeline.txt
"""
import ctypes as C
import os
import sys

import numpy as np

os.environ["M3S_GN_TIMELINE"] = "1"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
from mast3r_slam_b200 import _lib, kernels as K, synth  # noqa: E402


def main():
    # the stamps describe the dense chain over the whole system: keep the sparse front out unless asked for
    os.environ.setdefault("M3S_GN_SPARSE", "0")
    dev = torch.device("cuda", 0)
    g = synth.make_graph(7, 200, 103, 48, 64, num_pts=512)
    plan = K.GNPlan(g["ii"], g["jj"], 200, 512, 1, dev)
    params = K.make_params(_lib.GN_POINTS, 1, 1, 0.0, 0.0, 1.5, 0.01)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for rep in range(4):
        _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))
        _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid), p(Q),
                                         0, plan.E, None))
        torch.cuda.synchronize()
        e0.record()
        _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), None))
        e1.record()
        torch.cuda.synchronize()
    print(f"solve_update (assemble + factor + back-substitution + retract): {e0.elapsed_time(e1) * 1e3:.1f} us")
    n = C.c_int64(0)
    ptr = lib.m3s_gn_plan_timeline(plan.handle, C.byref(n))
    assert ptr and n.value > 0, "timeline disabled"
    iface = {"shape": (int(n.value),), "typestr": "<i8", "data": (int(ptr), False), "version": 3, "strides": None}
    holder = type("_B", (), {"__cuda_array_interface__": iface})()
    raw = torch.as_tensor(holder, device=dev).cpu().numpy().astype(np.float64)
    nblk = raw.size // 24
    tl = raw[: 8 * nblk].reshape(4, nblk, 2)
    ph = raw[8 * nblk:].reshape(2, nblk, 8)
    t0 = tl[tl > 0].min()
    tl = (tl - t0) * 1e-3  # us
    ph = np.where(ph > 0, (ph - t0) * 1e-3, np.nan)
    path = int(lib.m3s_gn_plan_factor_path(plan.handle))
    if path == 2:
        # dataflow factorisation: stamps of the diagonal CTAs only
        #   tl[0, k]   = elimination of diagonal block k: [start, flags published]
        #   ph[0, k, 0] = M_k computed; ph[0, k, 3..6] = CTA d = k/2 + 1 consuming M_k: wait start, M_k seen,
        #   panel computed, its tiles up to date
        d = tl[0]
        print("block  elim[start  M done  published]   next CTA [wait  seen  panel  tiles]")
        for kb in list(range(0, 8)) + list(range(20, 24)) + list(range(nblk - 4, nblk)):
            print(f"{kb:4d}   {d[kb, 0]:8.1f} {ph[0, kb, 0]:8.1f} {d[kb, 1]:8.1f}      "
                  + " ".join(f"{ph[0, kb, i]:8.1f}" for i in (3, 4, 5, 6)))
        nt = (nblk + 1) // 2
        print("SM of the diagonal CTAs:", [int(v) for v in raw[16 * nblk: 16 * nblk + nt]])
        sm_of = raw[16 * nblk + 120: 16 * nblk + 352].astype(int) - 1
        sm_of = sm_of[sm_of >= 0]
        if sm_of.size:
            cnt = np.bincount(sm_of, minlength=160)
            dsm = [int(v) for v in raw[16 * nblk: 16 * nblk + nt]]
            print("CTAs resident on the SMs of the diagonal CTAs:", [int(cnt[v]) for v in dsm],
                  "| SMs with 0/1/2 CTAs:", [int((cnt[:148] == k).sum()) for k in (0, 1, 2)])
        dbg = raw[16 * nblk + 64: 16 * nblk + 64 + 49]
        if dbg[48] > 0:
            for nm, off in (("chain", 0), ("inverse", 16), ("helper", 32)):
                v = dbg[off:off + 16] - dbg[48]
                print(f"first elimination of diagonal CTA 3 (block 6), clk since entry, {nm:8s}:", " ".join(f"{int(x):5d}" if x > 0 else "    -" for x in v))
        el = ph[0, :, 0] - d[:, 0]
        pub = d[:, 1] - ph[0, :, 0]
        hop = ph[0, :, 4] - d[:, 1]
        pan = ph[0, :, 5] - ph[0, :, 4]
        upd = ph[0, :, 6] - ph[0, :, 5]
        print("means (us): elim32 %.2f | stores + publish %.2f | flag hop %.2f | panel %.2f | tile updates %.2f"
              % (np.nanmean(el), np.nanmean(pub), np.nanmean(hop[1::2]), np.nanmean(pan), np.nanmean(upd)))
        print(f"factorisation: {d[nblk - 1, 1] - d[0, 0]:.1f} us = {(d[nblk - 1, 1] - d[0, 0]) / nblk:.2f} us per block step")
        b = tl[3]
        st = b[:, 0][b[:, 0] > 0]
        print(f"back-substitution: first CTA start {st.min():.1f} us, block 0 published at {b[0, 1]:.1f} us "
              f"({b[0, 1] - st.min():.1f} us for {nblk} blocks)")
        return
    names = ["diag", "tile-solve", "tile-upd", "backsolve"]
    for k, nm in enumerate(names):
        d = tl[k]
        ok = d[:, 1] > 0
        dur = (d[ok, 1] - d[ok, 0])
        print(f"{nm:9s}: n={ok.sum():3d} first start {d[ok, 0].min():8.1f} us, last end {d[ok, 1].max():8.1f} us, "
              f"duration mean {dur.mean():6.2f} us (min {dur.min():.2f}, max {dur.max():.2f})")
    print("step  diag[start end]      panel[start end]     syrk[start end]")
    for kb in list(range(0, 6)) + list(range(20, 23)) + list(range(nblk - 3, nblk)):
        print(f"{kb:4d}  " + "   ".join(f"{tl[k, kb, 0]:8.1f} {tl[k, kb, 1]:8.1f}" for k in range(3)))
    # phases inside the diag CTA (block kb) and inside update tile (0,0) of step kb
    d = ph[0]
    ks = [k for k in range(1, nblk) if not np.isnan(d[k, :4]).any()]
    if ks:
        seg = np.array([[d[k, 0] - tl[0, k, 0], d[k, 1] - d[k, 0], d[k, 2] - d[k, 1], d[k, 3] - d[k, 2],
                         tl[0, k, 1] - d[k, 3]] for k in ks])
        print("diag CTA phases, mean us over blocks 1..: load %.2f | B-solve %.2f | syrk %.2f | chol32 %.2f | store(+inv) %.2f"
              % tuple(seg.mean(0)))
    u = ph[1]
    ks = [k for k in range(nblk) if not np.isnan(u[k, :2]).any()]
    if ks:
        seg = np.array([[u[k, 0] - tl[1, k, 0], tl[1, k, 1] - u[k, 0], u[k, 1] - tl[2, k, 0], tl[2, k, 1] - u[k, 1]]
                        for k in ks])
        print("tile (0,0) phases, mean us: load %.2f | panel solve %.2f | rank-32 products %.2f | store %.2f"
              % tuple(seg.mean(0)))
    per_step = (tl[1, nblk - 1, 1] - tl[0, 0, 0]) / nblk
    print(f"factorisation: {tl[1, nblk - 1, 1] - tl[0, 0, 0]:.1f} us = {per_step:.2f} us per block step")
    b = tl[3]
    print(f"back-substitution: first CTA start {b[:, 0].min():.1f} us, block 0 published at {b[0, 1]:.1f} us "
          f"({b[0, 1] - b[:, 0].min():.1f} us, {(b[0, 1] - b[nblk - 1, 1]) / max(nblk - 1, 1):.2f} us per block)")


if __name__ == "__main__":
    main()
