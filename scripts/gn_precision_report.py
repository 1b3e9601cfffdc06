#!/usr/bin/env python
"""Pose-update error of the two per-point arithmetics of the GN accumulate pass (f64 / mixed)
against the CPU oracle, on the parity graphs of tests/test_gpu_gn.py.  Prints one line per case:
error relative to the size of the update (bar: 1e-5).  Run on the GPU box:
    python scripts/gn_precision_report.py > gpurun_out/gn_precision.txt
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from mast3r_slam_b200 import kernels as K, synth  # noqa: E402
from oracle import oracle  # noqa: E402  (checker only)


def upd_err(out, ref, Twc):
    out, ref, Twc = (np.asarray(x, np.float64) for x in (out, ref, Twc))
    upd = np.abs(ref - Twc).max()
    return np.abs(out - ref).max() / max(upd, 1e-30), upd


def gargs(g):
    return [g[k] for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q")]


def main():
    oracle.build()
    cases = [(10, 3, 2048, "rays", 10), (40, 12, 600, "points", 10), (12, 4, 1500, "calib", 10),
             (8, 3, 2001, "rays", 1), (8, 3, 2001, "points", 1), (8, 3, 2001, "calib", 1),
             (200, 103, 512, "points", 3), (10, 3, 12288, "rays", 10), (10, 3, 12288, "calib", 10)]
    import torch
    dev = torch.device("cuda", 0)
    print("err = max |poses - oracle (f32 output)| / max |update|;  d64 = max |fp64 pose state (mixed) - "
          "fp64 pose state (f64)| / max |update|  (no float32 output rounding)")
    print(f"{'graph':34s} {'mode':7s} it {'update':>9s} {'err f64':>10s} {'err mixed':>10s} {'d64':>10s}")
    for nkf, loops, npts, mode, it in cases:
        g = synth.make_graph(10 + nkf, nkf, loops, 96, 128, num_pts=npts)
        errs, st64 = [], []
        for prec in ("f64", "mixed"):
            K.set_gn_precision(prec)
            if mode == "calib":
                out = K.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"],
                                           g["valid"], g["Q"], g["img_size"], sigma_depth=10.0, max_iter=it)
            else:
                out = getattr(K, "gauss_newton_" + mode)(*gargs(g), max_iter=it)
            if prec == "f64":
                if mode == "calib":
                    ref = oracle.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"],
                                                    g["idx"], g["valid"], g["Q"], g["img_size"],
                                                    sigma_depth=10.0, max_iter=it)
                else:
                    ref = getattr(oracle, "gauss_newton_" + mode)(*gargs(g), max_iter=it)
            e, upd = upd_err(out, ref, g["Twc"])
            errs.append(e)
            plan = K.get_plan(g["ii"], g["jj"], nkf, npts, 1, dev)
            torch.cuda.synchronize()
            st64.append(plan.poses().cpu().numpy().copy())
        print(f"K={nkf:<3d} E={len(g['ii']):<4d} N={npts:<6d}          {mode:7s} {it:2d} {upd:9.2e} "
              f"{errs[0]:10.2e} {errs[1]:10.2e} {np.abs(st64[1] - st64[0]).max() / upd:10.2e}")
    K.set_gn_precision(None)


if __name__ == "__main__":
    main()
