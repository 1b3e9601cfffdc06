#!/usr/bin/env python
"""Does iter_proj (LSU / fp64 bound) overlap with refine_matches (FMA-issue bound) when they run
concurrently on two streams over different halves of the batch?  Prints sequential vs concurrent time."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mast3r_slam_b200 import kernels, synth

dev = torch.device("cuda", 0)
B, H, W, D = 32, 384, 512, 24
pairs = synth.make_pairs_torch(range(B), dev, H, W, D, "dyadic")
rays, pts, p0 = kernels.prep_for_iter_proj(pairs["X11"], pairs["X21"], None)
D11, D21 = pairs["D11"], pairs["D21"].reshape(B, H * W, D)
h = B // 2
s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
p_all, _ = kernels.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
torch.cuda.synchronize()

def seq():
    kernels.iter_proj(rays[:h], pts[:h], p0[:h], 10, 1e-8, 1e-6)
    kernels.refine_matches(D11[h:], D21[h:], p_all[h:], 3, 0)

def conc():
    with torch.cuda.stream(s1):
        kernels.iter_proj(rays[:h], pts[:h], p0[:h], 10, 1e-8, 1e-6)
    with torch.cuda.stream(s2):
        kernels.refine_matches(D11[h:], D21[h:], p_all[h:], 3, 0)

for name, fn in (("sequential", seq), ("concurrent", conc), ("sequential", seq), ("concurrent", conc)):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_stream(torch.cuda.current_stream()); s2.wait_stream(torch.cuda.current_stream())
    for _ in range(10):
        fn()
    torch.cuda.current_stream().wait_stream(s1); torch.cuda.current_stream().wait_stream(s2)
    e1.record()
    torch.cuda.synchronize()
    print(f"{name}: {e0.elapsed_time(e1) / 10:.4f} ms per (iter_proj on {h} pairs + refine on {h} pairs)")
