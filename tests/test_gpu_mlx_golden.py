"""GPU parity tests against golden vectors produced by EXECUTING the reference's MLX-only host code
(matching.py, image.py, geometry.py, global_opt.py, tracker.py) unmodified on a numpy-backed
``mlx.core`` stand-in (tests/golden/gen_golden_mlx.py) -- SURVEY.md s8 rows a1, a4, f1, f2, f3.
Every call goes through the C ABI (libm3s_b200.so); no oracle in between."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import config, global_opt, kernels, matching, tracker
    assert kernels.is_available()
    return kernels, matching, global_opt, tracker, config


def test_prep_for_iter_proj_vs_reference(mods, golden):
    K = mods[0]
    g = golden("mlx_match")
    for tag, init in (("none", None), ("init", g["idx_init"])):
        rays, pts, p0 = K.prep_for_iter_proj(g["X11"], g["X21"], init)
        assert np.array_equal(rays, g[f"rays_{tag}"]) and rays.dtype == np.float32
        assert np.array_equal(pts, g[f"pts_{tag}"])
        assert np.array_equal(p0, g[f"p_init_{tag}"])
    # image.img_gradient (image.py:9-34): channels 3-5 / 6-8 of the ray image, [B,C,H,W] there
    rays, _, _ = K.prep_for_iter_proj(g["X11"], g["X21"], None)
    unit = np.transpose(rays[..., :3], (0, 3, 1, 2))
    gx = np.zeros_like(unit); gx[:, :, :, 1:-1] = (unit[:, :, :, 2:] - unit[:, :, :, :-2]) / np.float32(2)
    assert np.array_equal(np.transpose(rays[..., 3:6], (0, 3, 1, 2)), gx)


def test_match_iterative_proj_whole_call_vs_reference(mods, golden):
    _, M, _, _, config = mods
    g = golden("mlx_match")
    a = (g["X11"], g["X21"], g["D11"], g["D21"])
    config.set_config({"matching": {"use_simple": False}})
    for tag, init in (("none", None), ("init", g["idx_init"])):
        idx, valid = M.match_iterative_proj(*a, init)
        assert idx.dtype == np.int32 and valid.dtype == np.bool_ and valid.shape == g[f"mip_valid_{tag}"].shape
        assert np.array_equal(idx, g[f"mip_idx_{tag}"]), tag
        assert np.array_equal(valid, g[f"mip_valid_{tag}"]), tag
        idx, valid = M.match(*a, init)  # the config switch (matching.py:33-38)
        assert np.array_equal(idx, g[f"mip_idx_{tag}"]) and np.array_equal(valid, g[f"mip_valid_{tag}"])
    config.set_config({"matching": {"use_simple": False, "refine_radius": 2, "dist_thresh": 0.3, "max_iter": 4}})
    idx, valid = M.match_iterative_proj(*a)
    assert np.array_equal(idx, g["mip_idx_r2"]) and np.array_equal(valid, g["mip_valid_r2"])
    config.set_config({"matching": {"use_simple": False, "use_refine": False}})
    idx, valid = M.match_iterative_proj(*a)
    assert np.array_equal(idx, g["mip_idx_norefine"]) and np.array_equal(valid, g["mip_valid_norefine"])


def test_match_default_route_vs_reference(mods, golden):
    """matching.match under the reference's default configuration (use_simple=True)."""
    import torch
    K, M, _, _, config = mods
    g = golden("mlx_match")
    a = (g["X11"], g["X21"], g["D11"], g["D21"])
    for tag, init in (("none", None), ("init", g["idx_init"])):
        idx, valid = M.match(*a, init)
        assert idx.dtype == np.int32 and np.array_equal(idx, g[f"simple_idx_{tag}"])
        assert valid.shape == g[f"simple_valid_{tag}"].shape and np.array_equal(valid, g[f"simple_valid_{tag}"])
    config.set_config({"matching": {"dist_thresh": 0.25}})
    _, valid = M.match(*a, g["idx_init"])
    assert np.array_equal(valid, g["simple_valid_thr025"])
    # device tensors in -> device tensors out; out-of-range indices are invalid, never read
    t = [torch.from_numpy(x).cuda() for x in a]
    bad = torch.from_numpy(g["idx_init"].copy()).cuda()
    bad[0, 0], bad[0, 1] = -1, g["idx_init"].shape[1]
    idx, valid = M.match(*t, bad)
    assert idx.is_cuda and not bool(valid[0, 0, 0]) and not bool(valid[0, 1, 0])
    assert np.array_equal(valid.cpu().numpy()[0, 2:], g["simple_valid_thr025"][0, 2:])


def test_constrain_points_to_ray_vs_reference(mods, golden):
    K = mods[0]
    g = golden("mlx_graph")
    h, w = (int(v) for v in g["hw"])
    out = K.constrain_points_to_ray((h, w), g["Xs"], g["K3"])
    assert np.array_equal(out, g["ray_Xs"])


def test_factor_graph_vs_reference(mods, golden):
    """add_factors (two batches + a refused relocalisation), the stored factors, the two-way edge
    set and solve_GN_calib, against the reference's FactorGraph run on real Keyframes / Frame objects."""
    _, _, G, _, config = mods
    g = golden("mlx_graph")
    h, w = (int(v) for v in g["hw"])
    n = h * w
    cfg = dict(config.get_config()["local_opt"])
    cfg.update(max_iters=4, sigma_depth=10.0)
    store = G.KeyframeStore(n, capacity=2)
    for k in range(len(g["Twc"])):
        nu = int(g["n_updates"][k])
        store.append(g["Xs"][k], g["Cs"][k] * nu, g["Twc"][k], nu)
    fg = G.FactorGraph(store, K=g["K3"], img_size=(h, w), cfg=cfg, capacity=1)
    E = len(g["ii_u"])
    half = E // 2
    mo = lambda s: (g["idx_i2j"][s], g["idx_j2i"][s], g["valid_j"][s], g["valid_i"][s], g["Qii"][s],  # noqa: E731
                    g["Qjj"][s], g["Qji"][s], g["Qij"][s])
    mmf = float(g["min_match_frac"])
    r1 = fg.add_factors(g["ii_u"][:half].tolist(), g["jj_u"][:half].tolist(), mmf, mo(slice(0, half)))
    r2 = fg.add_factors(g["ii_u"][half:].tolist(), g["jj_u"][half:].tolist(), mmf, mo(slice(half, E)))
    bad = int(np.flatnonzero((g["valid_j"].reshape(E, -1).sum(1) == 3))[0])
    r3 = fg.add_factors([int(g["ii_u"][bad])], [int(g["jj_u"][bad])], mmf, mo(slice(bad, bad + 1)), is_reloc=True)
    assert [r1, r2, r3] == [bool(v) for v in g["add_ret"]]
    for name in ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii"):
        got = getattr(fg, name).cpu().numpy()
        ref = g["fg_" + name]
        assert np.array_equal(got.reshape(ref.shape), ref), name
    # two-way edges: ours are pair-interleaved views; as a set of rows they are the reference's concat
    ii2, jj2, idx2, valid2, Q2 = (t.cpu().numpy() for t in fg._prep_two_way_edges())
    e = len(fg)
    order = np.r_[np.arange(0, 2 * e, 2), np.arange(1, 2 * e, 2)]
    assert np.array_equal(ii2[order], g["two_way_ii"]) and np.array_equal(jj2[order], g["two_way_jj"])
    assert np.array_equal(idx2[order], g["two_way_idx"])
    assert np.array_equal(valid2[order][..., None], g["two_way_valid"])
    assert np.array_equal(Q2[order][..., None], g["two_way_Q"])
    assert np.array_equal(fg.get_unique_kf_idx().cpu().numpy(), g["unique_kf"])
    fg.solve_GN_calib()
    out = store.T_WC[:len(store)].cpu().numpy().astype(np.float64)
    ref = g["poses_after_calib"].astype(np.float64)
    upd = np.abs(ref - g["Twc"]).max()
    assert upd > 1e-3
    assert np.abs(out - ref).max() <= 1e-5 * upd + 1.2e-7, np.abs(out - ref).max() / upd


TRACK_REL = 2e-3  # the reference iterates in float32 (MLX): its own noise, sized in test_oracle_mlx_golden


@pytest.mark.parametrize("tag", ["a", "b"])
def test_tracker_vs_reference(mods, golden, tag):
    _, _, _, T, config = mods
    g = golden("mlx_track")
    h, w = (int(v) for v in g[f"{tag}_hw"])
    Qk, v_opt, v_kf, _ = T.track_gate(g[f"{tag}_Cf"], g[f"{tag}_Ck"], g[f"{tag}_Qff"], g[f"{tag}_Qkf"],
                                      g[f"{tag}_idx_f2k"], g[f"{tag}_valid_match"])
    assert np.array_equal(Qk, g[f"{tag}_ray_Qk"]) and np.array_equal(v_opt, g[f"{tag}_ray_valid_opt"])
    start = g[f"{tag}_T_WCf"].astype(np.float64)
    for mode in ("ray", "calib"):
        t = f"{tag}_{mode}"
        args = (g[f"{t}_Xf_in"], g[f"{t}_Xk_in"], g[f"{tag}_T_WCf"], g[f"{tag}_T_WCk"], g[f"{t}_Qk"],
                g[f"{t}_valid_opt"])
        if mode == "ray":
            Tw, Tr = T.opt_pose_ray_dist_sim3(*args)
        else:
            Tw, Tr = T.opt_pose_calib_sim3(*args, g[f"{tag}_K4"], (h, w))
        ref_w, ref_r = g[f"{t}_T_WCf"].astype(np.float64), g[f"{t}_T_CkCf"].astype(np.float64)
        upd = np.abs(ref_w - start).max()
        assert np.abs(Tw - ref_w).max() <= TRACK_REL * upd, (mode, np.abs(Tw - ref_w).max() / upd)
        assert np.abs(Tr - ref_r).max() <= TRACK_REL * upd
        # the whole track() (tracker.py:51-180) on arrays
        config.set_config({"tracking": {"use_calib": mode == "calib"}})
        trk = T.FrameTracker()
        res = trk.track_arrays(g[f"{tag}_idx_f2k"], g[f"{tag}_valid_match"], g[f"{tag}_Xf"], g[f"{tag}_Cf"],
                               g[f"{tag}_Qff"], g[f"{tag}_Xk"], g[f"{tag}_Ck"], g[f"{tag}_Qkf"],
                               g[f"{tag}_T_WCf"], g[f"{tag}_T_WCk"], K=g[f"{tag}_K4"], img_size=(h, w))
        assert res["new_kf"] == bool(g[f"{t}_track_new_kf"]) and res["try_reloc"] == bool(g[f"{t}_track_try_reloc"])
        assert np.abs(res["T_WCf"] - g[f"{t}_track_T_WCf"].astype(np.float64)).max() <= TRACK_REL * upd
