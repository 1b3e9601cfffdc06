"""Golden vectors of the Metal-semantics mode: outputs of the reference's
UNMODIFIED Metal compute shaders, compiled as C++ where they lie under /root/reference and run one
"thread" at a time (oracle/metal_ref.cpp + oracle/metal_shim/metal_stdlib), driven by the numpy
restatement of the reference's Metal runners (oracle/metal_oracle.py).  Runs only in the build
container; the small .npz fixtures are committed.

    python tests/golden/gen_golden_metal.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

import mast3r_slam_b200.synth as synth  # noqa: E402
from oracle import metal_oracle as mo  # noqa: E402
from oracle import oracle  # noqa: E402  (prep step only)

assert os.path.isdir(mo.REF_SHADERS), "needs the reference's shader sources"
mo.build()


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **kw)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def gen_match():
    out = {}
    # incoherent fixture (value distributions of benchmark_all_kernels.py:67-76), starts outside the image
    rays, pts, p0 = synth.reference_like_iter_proj(21, 2, 40, 56, 1200)
    p0[0, :7] = np.array([[-3.5, 4], [60.2, 10], [5, -2.2], [7, 44.9], [55.0, 39.0], [0, 0],
                          [54.999, 38.999]], np.float32)
    for it in (1, 10):
        p, v = mo.iter_proj(rays, pts, p0, it, 1e-8, 1e-6)
        out[f"rnd_p_{it}"], out[f"rnd_valid_{it}"] = p, v
    out.update(rnd_rays=rays, rnd_pts=pts, rnd_p0=p0)
    # coherent scene; a loose threshold makes the PER-POINT exit fire at different iterations
    pair = synth.make_pair(5, 48, 64)
    rays, pts, p0 = oracle.prep_for_iter_proj_numpy(pair["X11"][None], pair["X21"][None], None)
    for tag, it, th in (("a", 10, 1e-6), ("b", 10, 2e-2), ("c", 3, 1e-6)):
        p, v = mo.iter_proj(rays, pts, p0, it, 1e-8, th)
        out[f"scn_p_{tag}"], out[f"scn_valid_{tag}"] = p, v
    out.update(scn_rays=rays, scn_pts=pts, scn_p0=p0)
    # refine cascade: unit-norm descriptors (sequential fp32 sums), dilation_max 1, 2, 5, radius 1..3
    D11, D21, p1 = synth.reference_like_refine(22, 2, 36, 44, 24, 900)
    p1[0, :4] = np.array([[0, 0], [43, 35], [-2, 5], [50, 40]], np.int32)   # border / outside centres
    for r, dm in ((3, 1), (3, 2), (2, 5), (1, 0)):
        out[f"ref_p1_r{r}_d{dm}"] = mo.refine_matches(D11, D21, p1, r, dm)
    out.update(ref_D11=D11, ref_D21=D21, ref_p1=p1)
    D11, D21, p1 = synth.reference_like_refine(23, 1, 20, 28, 16, 300)
    out.update(ref16_D11=D11, ref16_D21=D21, ref16_p1=p1, ref16_out=mo.refine_matches(D11, D21, p1, 3, 2))
    D11, D21, p1 = synth.reference_like_refine(24, 1, 12, 14, 10, 100)    # D not a multiple of 4
    out.update(ref10_D11=D11, ref10_D21=D21, ref10_p1=p1, ref10_out=mo.refine_matches(D11, D21, p1, 2, 3))
    save("metal_match", **out)


def gen_gn():
    g = synth.make_graph(31, 6, 2, 32, 40, num_pts=700)
    a = (g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
    out = {k: g[k] for k in ("Twc", "Twc_gt", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q", "K4")}
    out["img_size"] = np.array(g["img_size"])
    for it in (1, 4):
        out[f"rays_{it}"] = mo.gauss_newton_rays(*a, max_iter=it)
        out[f"points_{it}"] = mo.gauss_newton_points(*a, max_iter=it)
        out[f"calib_{it}"] = mo.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"],
                                                   g["valid"], g["Q"], g["img_size"], sigma_depth=10.0, max_iter=it)
    dbg = {}
    mo.gauss_newton_rays(*a, max_iter=1, debug=dbg)
    out["rays_H"], out["rays_g"] = dbg["H"], dbg["g"]
    # what the numpy path gives on the same input (the behavioural gap the mode exists for)
    out["numpy_rays_4"] = oracle.gauss_newton_rays(*a, max_iter=4)
    save("metal_gn", **out)


if __name__ == "__main__":
    gen_match()
    gen_gn()
