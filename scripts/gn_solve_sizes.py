"""solve_update time of the sparse front against the dense paths on trajectories of several lengths
(window-2 chain + loop closures, 256 points per keyframe: the solve does not depend on the point count).
Usage (GPU box): python scripts/gn_solve_sizes.py"""
import ctypes as C
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from mast3r_slam_b200 import _lib, kernels as K, synth  # noqa: E402


def run(num_kf, loops, env):
    for k in ("M3S_GN_CHOL", "M3S_GN_SPARSE"):
        os.environ.pop(k, None)
    os.environ.update(env)
    dev = torch.device("cuda", 0)
    g = synth.make_graph(31, num_kf, loops, 24, 32, num_pts=256)
    plan = K.GNPlan(g["ii"], g["jj"], num_kf, 256, 1, dev)
    params = K.make_params(_lib.GN_POINTS, 1, 1, 0.0, 0.0, 1.5, 0.01)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    best = 1e9
    for rep in range(6):
        _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))
        _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid), p(Q),
                                         0, plan.E, None))
        torch.cuda.synchronize()
        e0.record()
        _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), None))
        e1.record()
        torch.cuda.synchronize()
        if rep >= 2:
            best = min(best, e0.elapsed_time(e1) * 1e3)
    return best, plan.factor_path(), plan.sparse_info(), plan.status()


if __name__ == "__main__":
    for num_kf, loops in ((100, 30), (200, 103), (200, 20), (500, 30), (1000, 60)):
        a = run(num_kf, loops, {})
        b = run(num_kf, loops, {"M3S_GN_SPARSE": "0"})
        print(f"{num_kf:5d} KF {loops:4d} loops: default path {a[1]} {a[0]:8.1f} us {a[2]} | dense path {b[1]} {b[0]:8.1f} us"
              f" | failed {a[3]['failed']} {b[3]['failed']}")
