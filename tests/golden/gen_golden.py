"""Generate golden input/output vectors by RUNNING THE REFERENCE's own numpy ("CPU path")
implementation, imported unmodified from /root/reference/src.  Runs only in the build container
(the reference does not travel to the GPU box); the resulting small .npz fixtures are committed.

    python tests/golden/gen_golden.py

The reference has no tests or stored fixtures, so these vectors -- outputs of the
reference itself on seeded inputs -- are what pins the oracle (tests/test_oracle_golden.py) and,
through it, the CUDA path.
"""
import os
import sys
import warnings

import numpy as np

sys.dont_write_bytecode = True  # /root/reference is read-only
REF_SRC = "/root/reference/src"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, REF_SRC)
sys.path.insert(0, ROOT)

from mlx_mast3r_slam.backends.mpsgraph import kernels as ref  # noqa: E402

import mast3r_slam_b200.synth as synth  # noqa: E402
from oracle import oracle  # noqa: E402  (only for the MLX-only host steps restated in numpy)

assert not ref.is_available(), "expected the numpy fallbacks (no Metal here)"
warnings.filterwarnings("ignore")


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **kw)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def gen_iter_proj():
    # (a) incoherent worst case, value distributions of benchmark_all_kernels.py:67-76
    rays, pts, p0 = synth.reference_like_iter_proj(11, 2, 40, 56, 1500)
    # a few starts outside the image to exercise clamp/valid
    p0[0, :8] = np.array([[-3.5, 4], [60.2, 10], [5, -2.2], [7, 44.9], [55.0, 39.0], [0, 0],
                          [54.999, 38.999], [1e6, -1e6]], np.float32)
    p, v = ref._iter_proj_numpy(rays, pts, p0.copy(), 10, 1e-8, 1e-6)
    save("iter_proj_random", rays=rays, pts=pts, p_init=p0, p=p, valid=v)
    # (b) coherent scene through the (restated) prep step, 3 and 10 iterations
    pair = synth.make_pair(3, 48, 64)
    rays, pts, p0 = oracle.prep_for_iter_proj_numpy(pair["X11"][None], pair["X21"][None], None)
    out = {}
    for it in (1, 3, 10):
        p, v = ref._iter_proj_numpy(rays, pts, p0.copy(), it, 1e-8, 1e-6)
        out[f"p_{it}"] = p
        out[f"valid_{it}"] = v
    save("iter_proj_scene", X11=pair["X11"], X21=pair["X21"], rays=rays, pts=pts, p_init=p0, **out)
    # (c) global early exit: view 2 == view 1 -> delta == 0 at the first iteration; and a large
    # threshold that stops after the 2nd iteration (kernels.py:239-241)
    rays, pts, p0 = oracle.prep_for_iter_proj_numpy(pair["X11"][None], pair["X11"][None], None)
    p_same, v_same = ref._iter_proj_numpy(rays, pts, p0.copy(), 10, 1e-8, 1e-6)
    rays2, pts2, p02 = oracle.prep_for_iter_proj_numpy(pair["X11"][None], pair["X21"][None], None)
    # pick a threshold between the 2nd- and 3rd-iteration max step on interior points only
    sel = np.zeros((48, 64), bool)
    sel[8:40, 8:56] = True
    sel = sel.reshape(-1)
    pts_i, p0_i = pts2[:, sel], p02[:, sel]
    prev = p0_i.copy()
    steps = []
    for it in range(1, 6):
        pi, _ = ref._iter_proj_numpy(rays2, pts_i, p0_i.copy(), it, 1e-8, 0.0)
        steps.append(float(np.sqrt(((pi - prev) ** 2).sum(-1)).max()))
        prev = pi
    thr = float(np.sqrt(steps[1] * steps[2]))
    p_thr, v_thr = ref._iter_proj_numpy(rays2, pts_i, p0_i.copy(), 10, 1e-8, thr)
    save("iter_proj_break", rays_same=rays, pts_same=pts, p0_same=p0, p_same=p_same,
         valid_same=v_same, rays=rays2, pts=pts_i, p_init=p0_i, thresh=np.float64(thr),
         steps=np.array(steps), p=p_thr, valid=v_thr)


def gen_refine():
    # (a) dyadic descriptors from the coherent scene: exact dot products under any order
    pair = synth.make_pair(5, 40, 56, 24, "dyadic")
    h, w = 40, 56
    rng = np.random.default_rng(7)
    n = 700
    sel = rng.choice(h * w, n, replace=False)
    uv = pair["flow_uv"].reshape(-1, 2)[sel]
    p1 = np.trunc(uv + rng.uniform(-2, 2, uv.shape)).astype(np.int32)
    # border / out-of-image centres (window partly or wholly outside)
    p1[:10] = np.array([[0, 0], [w - 1, h - 1], [-2, 5], [w + 1, 3], [4, -3], [9, h + 2],
                        [-10, -10], [w + 9, h + 9], [1, h - 2], [w - 2, 1]], np.int32)
    D11 = pair["D11"][None]
    D21 = pair["D21"].reshape(-1, 24)[sel][None]
    out = {}
    for r, dil in ((3, 0), (3, 2), (1, 1), (2, 3)):
        out[f"r{r}_d{dil}"] = ref._refine_matches_numpy(D11, D21, p1[None], r, dil)
    save("refine_dyadic", D11=D11, D21=D21, p1=p1[None], **out)
    # (b) unit-norm gaussian descriptors (reference generator's distribution), D=32, b=2
    D11, D21, p1 = synth.reference_like_refine(13, 2, 24, 32, 32, 300)
    o = ref._refine_matches_numpy(D11, D21, p1, 3, 2)
    save("refine_unit", D11=D11, D21=D21, p1=p1, r3_d2=o)
    # (c) float p1 input (the API accepts any dtype and truncates, kernels.py:508,515)
    p1f = (p1.astype(np.float32) + 0.7)[:, :50]
    o = ref._refine_matches_numpy(D11, D21[:, :50], p1f, 2, 0)
    save("refine_floatp", p1f=p1f, r2_d0=o)


def gen_gn():
    # (a) the reference generator's distributions (random poses, permutation matches): 3 iterations
    Twc, Xs, Cs, ii, jj, idx, valid, Q = synth.reference_like_gn(42, 5, 200, 8)
    o_r = ref.gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, max_iter=3)
    o_p = ref.gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, max_iter=3)
    save("gn_random", Twc=Twc, Xs=Xs, Cs=Cs, ii=ii, jj=jj, idx=idx, valid=valid, Q=Q,
         rays_it3=o_r, points_it3=o_p,
         rays_it1=ref.gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, max_iter=1),
         points_it1=ref.gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, max_iter=1))
    # (b) physically consistent 5-KF graph, [...,1]-shaped Cs/valid/Q, pin=1 and pin=2
    g = synth.make_graph(1, 5, 1, 24, 32, num_pts=240)
    kw = dict(Twc=g["Twc"], Xs=g["Xs"], Cs=g["Cs"][..., None], ii=g["ii"], jj=g["jj"],
              idx_ii2jj=g["idx"], valid_match=g["valid"][..., None], Q=g["Q"][..., None])
    out = {}
    for it in (1, 2, 6):
        out[f"rays_it{it}"] = ref.gauss_newton_rays(**kw, max_iter=it)
        out[f"points_it{it}"] = ref.gauss_newton_points(**kw, max_iter=it)
    out["rays_pin2"] = ref.gauss_newton_rays(**kw, max_iter=2, pin=2)
    K3 = np.array([[g["K4"][0], 0, g["K4"][2]], [0, g["K4"][1], g["K4"][3]], [0, 0, 1]], np.float32)
    for it in (1, 2, 6):
        out[f"calib_it{it}"] = ref.gauss_newton_calib(
            g["Twc"], g["Xs"], g["Cs"], K3, g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
            g["img_size"], pixel_border=1, z_eps=0.0, sigma_pixel=1.0, sigma_depth=10.0,
            max_iter=it)
    out["calib_k4"] = ref.gauss_newton_calib(
        g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"],
        g["img_size"], pixel_border=0, z_eps=0.0, sigma_pixel=1.0, sigma_depth=0.1, max_iter=2)
    save("gn_graph", Twc=g["Twc"], Twc_gt=g["Twc_gt"], Xs=g["Xs"], Cs=g["Cs"], ii=g["ii"],
         jj=g["jj"], idx=g["idx"], valid=g["valid"], Q=g["Q"], K4=g["K4"], K3=K3,
         img_size=np.array(g["img_size"]), **out)
    # (c) early-outs: no edges, all KFs pinned, edges touching only pinned KFs
    e0 = ref.gauss_newton_rays(g["Twc"], g["Xs"], g["Cs"], g["ii"][:0], g["jj"][:0], g["idx"][:0],
                               g["valid"][:0], g["Q"][:0])
    pin_all = ref.gauss_newton_rays(**kw, max_iter=2, pin=5)
    save("gn_early", no_edges=e0, pin_all=pin_all)


if __name__ == "__main__":
    gen_iter_proj()
    gen_refine()
    gen_gn()
