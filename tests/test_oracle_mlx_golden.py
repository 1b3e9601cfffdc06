"""Pins the restatements of the reference's MLX-only host steps
against golden vectors produced by EXECUTING the unmodified reference modules on the numpy-backed
``mlx.core`` stand-in (tests/golden/gen_golden_mlx.py).  Integer / boolean outputs and the float32
element-wise steps are bit-exact; the float32 iterative pose solves are compared to the float64
restatement with the tolerance the reference's own precision allows."""
import numpy as np
import pytest

from oracle import oracle as orc


def test_prep_bit_exact_vs_reference(oracle, golden):
    g = golden("mlx_match")
    for tag, init in (("none", None), ("init", g["idx_init"])):
        for fn in (oracle.prep_for_iter_proj, oracle.prep_for_iter_proj_numpy):
            rays, pts, p0 = fn(g["X11"], g["X21"], init)
            assert np.array_equal(rays, g[f"rays_{tag}"]), (tag, fn.__name__)
            assert np.array_equal(pts, g[f"pts_{tag}"])
            assert np.array_equal(p0, g[f"p_init_{tag}"])
    # img_gradient (image.py:9-34) is channels 3..8 of the ray image
    unit = g["rays_none"][..., :3]
    rays, _, _ = oracle.prep_for_iter_proj(g["X11"], g["X21"], None)
    assert np.array_equal(rays[..., :3], unit)


def test_match_iterative_proj_whole_call_vs_reference(oracle, golden):
    g = golden("mlx_match")
    for tag, init in (("none", None), ("init", g["idx_init"])):
        idx, valid = oracle.match_iterative_proj(g["X11"], g["X21"], g["D11"], g["D21"], init)
        assert np.array_equal(idx, g[f"mip_idx_{tag}"]), tag
        assert np.array_equal(valid, g[f"mip_valid_{tag}"]), tag
    idx, valid = oracle.match_iterative_proj(g["X11"], g["X21"], g["D11"], g["D21"], None, max_iter=4,
                                             radius=2, dist_thresh=0.3)
    assert np.array_equal(idx, g["mip_idx_r2"]) and np.array_equal(valid, g["mip_valid_r2"])
    idx, valid = oracle.match_iterative_proj(g["X11"], g["X21"], g["D11"], g["D21"], None, radius=0)
    assert np.array_equal(idx, g["mip_idx_norefine"]) and np.array_equal(valid, g["mip_valid_norefine"])
    assert 50 < g["mip_valid_none"].sum() < g["mip_valid_none"].size  # the fixture is not degenerate


def test_match_simple_vs_reference(oracle, golden):
    g = golden("mlx_match")
    for tag, init in (("none", None), ("init", g["idx_init"])):
        idx, valid = oracle.match_simple(g["X11"], g["X21"], init)
        assert np.array_equal(idx, g[f"simple_idx_{tag}"]) and idx.dtype == np.int32
        assert np.array_equal(valid, g[f"simple_valid_{tag}"]) and valid.shape == g[f"simple_valid_{tag}"].shape
    _, valid = oracle.match_simple(g["X11"], g["X21"], g["idx_init"], dist_thresh=0.25)
    assert np.array_equal(valid, g["simple_valid_thr025"])
    assert g["simple_valid_init"].sum() < g["simple_valid_thr025"].sum() < g["simple_valid_init"].size
    assert g["mip_valid_r2"].sum() > g["mip_valid_none"].sum()


def test_constrain_points_to_ray_vs_reference(golden):
    g = golden("mlx_graph")
    h, w = (int(v) for v in g["hw"])
    out = orc.constrain_points_to_ray_numpy((h, w), g["Xs"], g["K3"])
    assert np.array_equal(out, g["ray_Xs"])


def test_add_factors_and_two_way_edges_vs_reference(golden):
    g = golden("mlx_graph")
    E = len(g["ii_u"])
    half = E // 2
    rows = []
    for lo, hi in ((0, half), (half, E)):
        s = slice(lo, hi)
        r = orc.add_factors_numpy(g["ii_u"][s], g["jj_u"][s],
                                  (g["idx_i2j"][s], g["idx_j2i"][s], g["valid_j"][s], g["valid_i"][s],
                                   g["Qii"][s], g["Qjj"][s], g["Qji"][s], g["Qij"][s]), 1.5, float(g["min_match_frac"]))
        assert r is not None
        rows.append(r)
    cat = [np.concatenate([a[k] for a in rows]) for k in range(8)]
    names = ("fg_ii", "fg_jj", "fg_idx_ii2jj", "fg_idx_jj2ii", "fg_valid_match_j", "fg_valid_match_i",
             "fg_Q_ii2jj", "fg_Q_jj2ii")
    for a, nm in zip(cat, names):
        assert np.array_equal(a, g[nm]), nm
    assert len(g["fg_ii"]) == E - 1  # the starved loop-closure pair was dropped
    assert list(g["add_ret"]) == [True, True, False]
    # two-way concat (global_opt.py:147-154)
    assert np.array_equal(np.concatenate([cat[0], cat[1]]), g["two_way_ii"])
    assert np.array_equal(np.concatenate([cat[2], cat[3]]), g["two_way_idx"])
    assert np.array_equal(np.concatenate([cat[6], cat[7]]), g["two_way_Q"])


def test_solve_gn_calib_through_factor_graph_vs_reference(oracle, golden):
    """FactorGraph.solve_GN_calib end to end (global_opt.py:212-268): Cs = C / N, points constrained
    to rays, img_size handed over as (H, W) -- the reference's order, which gauss_newton_calib
    unpacks as (width, height) (gauss_newton_calib.py:75)."""
    g = golden("mlx_graph")
    h, w = (int(v) for v in g["hw"])
    Xs = orc.constrain_points_to_ray_numpy((h, w), g["Xs"], g["K3"])
    out = oracle.gauss_newton_calib(g["Twc"], Xs, g["Cs"], g["K3"], g["two_way_ii"], g["two_way_jj"],
                                    g["two_way_idx"], g["two_way_valid"], g["two_way_Q"], (h, w),
                                    pixel_border=0, z_eps=0.0, sigma_pixel=1.0, sigma_depth=10.0,
                                    C_thresh=0.0, Q_thresh=1.5, max_iter=4, delta_thresh=1e-3, pin=1)
    ref = g["poses_after_calib"].astype(np.float64)
    upd = np.abs(ref - g["Twc"]).max()
    assert upd > 1e-3
    assert np.abs(out - ref).max() <= 1e-5 * upd + 1.2e-7
    # with the order a (W, H)-reading caller would expect the result differs: the deviation is real
    out_wh = oracle.gauss_newton_calib(g["Twc"], Xs, g["Cs"], g["K3"], g["two_way_ii"], g["two_way_jj"],
                                       g["two_way_idx"], g["two_way_valid"], g["two_way_Q"], (w, h),
                                       pixel_border=0, z_eps=0.0, sigma_pixel=1.0, sigma_depth=10.0,
                                       C_thresh=0.0, Q_thresh=1.5, max_iter=4, delta_thresh=1e-3, pin=1)
    assert np.abs(out_wh - ref).max() > 10 * (1e-5 * upd + 1.2e-7)


@pytest.mark.parametrize("tag", ["a", "b"])
def test_track_gate_and_inputs_vs_reference(golden, tag):
    g = golden("mlx_track")
    Qk, v_opt, v_kf, _ = orc.track_gate_numpy(g[f"{tag}_Cf"], g[f"{tag}_Ck"], g[f"{tag}_Qff"], g[f"{tag}_Qkf"],
                                              g[f"{tag}_idx_f2k"], g[f"{tag}_valid_match"])
    assert np.array_equal(Qk, g[f"{tag}_ray_Qk"])
    assert np.array_equal(v_opt, g[f"{tag}_ray_valid_opt"])
    assert np.array_equal(g[f"{tag}_Xf"][g[f"{tag}_idx_f2k"]], g[f"{tag}_ray_Xf_in"])
    # calibrated mode: points constrained to the pixel rays first (tracker.py:196-197)
    h, w = (int(v) for v in g[f"{tag}_hw"])
    Xk_c = orc.constrain_points_to_ray_numpy((h, w), g[f"{tag}_Xk"][None], g[f"{tag}_K4"])[0]
    assert np.array_equal(Xk_c, g[f"{tag}_calib_Xk_in"])


# The reference iterates in float32 (MLX); the restatement in float64 is the exact-math statement of
# the same formulas.  The float32 run of the restatement must reproduce the reference's float32
# output closely (same formulas, same precision, reduction order differs), and the float64 run must
# agree within the float32 noise of the reference itself.
@pytest.mark.parametrize("tag", ["a", "b"])
@pytest.mark.parametrize("mode", ["ray", "calib"])
def test_track_solves_vs_reference(golden, tag, mode):
    g = golden("mlx_track")
    t = f"{tag}_{mode}"
    h, w = (int(v) for v in g[f"{tag}_hw"])
    Xf, Xk = g[f"{t}_Xf_in"], g[f"{t}_Xk_in"]
    args = (Xf, Xk, g[f"{tag}_T_WCf"], g[f"{tag}_T_WCk"], g[f"{t}_Qk"], g[f"{t}_valid_opt"])
    res = {}
    for dt in (np.float32, np.float64):
        if mode == "ray":
            res[dt] = orc.opt_pose_ray_dist_sim3_numpy(*args, dtype=dt)
        else:
            res[dt] = orc.opt_pose_calib_sim3_numpy(*args, g[f"{tag}_K4"], (h, w), dtype=dt)
    ref_w, ref_r = g[f"{t}_T_WCf"].astype(np.float64), g[f"{t}_T_CkCf"].astype(np.float64)
    upd = np.abs(ref_w - g[f"{tag}_T_WCf"]).max()
    assert upd > 1e-3
    e32 = max(np.abs(res[np.float32][0] - ref_w).max(), np.abs(res[np.float32][1] - ref_r).max())
    e64 = max(np.abs(res[np.float64][0] - ref_w).max(), np.abs(res[np.float64][1] - ref_r).max())
    assert e32 <= 2e-3 * upd, (e32, upd)
    assert e64 <= 2e-3 * upd, (e64, upd)
    # the whole track() lands on the same pose as the solver entry point
    assert np.array_equal(g[f"{t}_track_T_WCf"], g[f"{t}_T_WCf"])
    assert not bool(g[f"{t}_track_try_reloc"])
