"""Where the tracker solve's time goes: device time of m3s_track_opt_pose for several point counts and
iteration caps (rel_error / delta_norm = 0 keep the loop running to max_iters)."""
import ctypes as C
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_b200 import _lib, synth, tracker  # noqa: E402

dev = torch.device("cuda", 0)
lib = _lib.load()
for hw in ((16, 32), (96, 128), (384, 512)):
    d = synth.make_track_pair(0, *hw)
    n = hw[0] * hw[1]
    t = lambda a, dt=torch.float32: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Xf, Xk, idx = t(d["Xf"]), t(d["Xk"]), t(d["idx_f2k"], torch.int32)
    Qk = t(np.sqrt(d["Qff"][d["idx_f2k"]] * d["Qkf"]))
    vv = t(d["valid_match"].reshape(-1).astype(np.uint8), torch.uint8)
    Tf, Tk = t(d["T_WCf"]), t(d["T_WCk"])
    ws = torch.empty(lib.m3s_track_workspace_bytes(n), dtype=torch.uint8, device=dev)
    o1, o2 = torch.empty(8, device=dev), torch.empty(8, device=dev)
    stats = torch.empty(4, dtype=torch.float64, device=dev)
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    st = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    for iters in (0, 1, 2, 4, 8):
        cfg = dict(tracker.get_config()["tracking"], max_iters=iters, rel_error=0.0, delta_norm=0.0)
        prm = tracker._params(_lib.TRACK_RAY_DIST, cfg)
        call = lambda: _lib.check(lib.m3s_track_opt_pose(  # noqa: E731
            C.byref(prm), p(Xf), p(idx), p(Xk), p(Qk), p(vv), n, p(Tf), p(Tk), p(o1), p(o2), p(stats),
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
        print(f"N={n:7d} max_iters={iters}: {1e3 * e0.elapsed_time(e1) / 50:7.1f} us  (iterations run {int(stats.cpu()[0])})")
