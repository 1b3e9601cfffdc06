"""GPU parity tests for the tracker single-pose solves: track_gate,
opt_pose_ray_dist_sim3, opt_pose_calib_sim3, FrameTracker.track_arrays -- all through the C ABI
(m3s_track_gate / m3s_track_opt_pose).  The checker is the numpy restatement of tracker.py /
geometry.py / liegroups in oracle/oracle.py (float64); the reference evaluates the same formulas in
MLX float32, so the tolerance is the north-star 1e-5 relative to the pose update, and the float32
run of the restatement is required to sit inside the same band (it sizes the reference's own noise)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-5  # on the pose update, as for the graph solvers


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import kernels, synth, tracker
    assert kernels.is_available()
    return tracker, synth, kernels


def _gated(oracle, d):
    Qk, v_opt, v_kf, counts = oracle.track_gate_numpy(d["Cf"], d["Ck"], d["Qff"], d["Qkf"], d["idx_f2k"],
                                                     d["valid_match"])
    return Qk, v_opt, v_kf, counts


def _check_pose(out, ref, start, what):
    upd = np.abs(ref - start.astype(np.float64)).max()
    err = np.abs(out.astype(np.float64) - ref).max()
    assert err <= REL * upd + 1.2e-7, (what, err, upd)  # + 1 float32 ulp of the final cast


@pytest.mark.parametrize("seed,hw", [(0, (96, 128)), (1, (48, 64)), (2, (120, 160))])
def test_track_gate_bit_exact(mods, oracle, seed, hw):
    trk, synth, _ = mods
    d = synth.make_track_pair(seed, *hw)
    d["Cf"][::7] = 0.0  # below C_conf = 0 is impossible (strict >): make some fail
    ref = _gated(oracle, d)
    out = trk.track_gate(d["Cf"], d["Ck"], d["Qff"], d["Qkf"], d["idx_f2k"], d["valid_match"])
    assert np.array_equal(out[0], ref[0]) and out[0].dtype == np.float32 and out[0].shape == ref[0].shape
    assert np.array_equal(out[1], ref[1]) and np.array_equal(out[2], ref[2])
    assert out[3] == ref[3] and ref[3][0] < ref[3][1]


@pytest.mark.parametrize("seed,hw", [(0, (96, 128)), (3, (64, 80)), (4, (192, 256))])
def test_ray_dist_matches_oracle(mods, oracle, seed, hw):
    trk, synth, _ = mods
    d = synth.make_track_pair(seed, *hw)
    Qk, v_opt, _, _ = _gated(oracle, d)
    Xf = d["Xf"][d["idx_f2k"]]
    Tw, Tr, info = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt)
    out_w, out_r, st = trk.opt_pose_ray_dist_sim3(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt,
                                                  return_stats=True)
    assert out_w.dtype == np.float32 and out_w.shape == (8,)
    assert st["iterations"] == info["iterations"]
    assert abs(st["cost"] - info["cost"]) <= 1e-9 * abs(info["cost"])
    _check_pose(out_w, Tw, d["T_WCf"], "T_WCf")
    _check_pose(out_r, Tr, d["T_WCf"], "T_CkCf")
    # the solve moves the pose towards the ground truth
    assert np.abs(out_w - d["T_WCf_gt"]).max() < 0.1 * np.abs(d["T_WCf"] - d["T_WCf_gt"]).max()
    # fused gather (idx_f2k passed to the kernel) is the same computation
    out_w2, out_r2 = trk.opt_pose_ray_dist_sim3(d["Xf"], d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt,
                                                idx_f2k=d["idx_f2k"])
    assert np.array_equal(out_w, out_w2) and np.array_equal(out_r, out_r2)
    # the reference's own precision (float32 restatement) lies within the same band
    Tw32, _, _ = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt,
                                                     dtype=np.float32)
    assert np.abs(Tw32 - Tw).max() <= 1e-3 * np.abs(Tw - d["T_WCf"]).max() + 1e-6


@pytest.mark.parametrize("seed,hw,border", [(0, (96, 128), 0), (5, (64, 80), 3)])
def test_calib_matches_oracle(mods, oracle, seed, hw, border):
    trk, synth, K_ = mods
    d = synth.make_track_pair(seed, *hw)
    Qk, v_opt, _, _ = _gated(oracle, d)
    n = hw[0] * hw[1]
    Xf_c = K_.constrain_points_to_ray(hw, d["Xf"][None], d["K4"])[0]
    Xk_c = K_.constrain_points_to_ray(hw, d["Xk"][None], d["K4"])[0]
    Xk_c[11, 2] = -0.5  # a measurement behind the camera is masked (tracker.py:207)
    cfg = dict(oracle.DEFAULT_TRACKING, pixel_border=border)
    Xf = Xf_c[d["idx_f2k"]]
    Tw, Tr, info = oracle.opt_pose_calib_sim3_numpy(Xf, Xk_c, d["T_WCf"], d["T_WCk"], Qk, v_opt, d["K4"],
                                                    hw, cfg)
    out_w, out_r, st = trk.opt_pose_calib_sim3(Xf, Xk_c, d["T_WCf"], d["T_WCk"], Qk, v_opt, d["K4"], hw,
                                               cfg, return_stats=True)
    assert st["iterations"] == info["iterations"] and n == len(Xk_c)
    assert abs(st["cost"] - info["cost"]) <= 1e-9 * abs(info["cost"])
    _check_pose(out_w, Tw, d["T_WCf"], "T_WCf")
    _check_pose(out_r, Tr, d["T_WCf"], "T_CkCf")
    K33 = np.array([[d["K4"][0], 0, d["K4"][2]], [0, d["K4"][1], d["K4"][3]], [0, 0, 1]], np.float32)
    out_w3, _ = trk.opt_pose_calib_sim3(Xf, Xk_c, d["T_WCf"], d["T_WCk"], Qk, v_opt, K33, hw, cfg)
    assert np.array_equal(out_w3, out_w)


def test_edge_cases(mods, oracle):
    trk, synth, _ = mods
    d = synth.make_track_pair(6, 48, 64)
    Qk, v_opt, _, _ = _gated(oracle, d)
    Xf = d["Xf"][d["idx_f2k"]]
    # max_iters = 0: T_WCk * (T_WCk^-1 * T_WCf), no update
    cfg0 = dict(oracle.DEFAULT_TRACKING, max_iters=0)
    Tw, Tr, _ = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt, cfg0)
    out_w, out_r = trk.opt_pose_ray_dist_sim3(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt, cfg0)
    assert np.abs(out_w - Tw).max() <= 1.2e-7 and np.abs(out_r - Tr).max() <= 1.2e-7
    # one iteration only, and no valid point at all (H = 1e-6 I, tau = 0 -> converged at once)
    cfg1 = dict(oracle.DEFAULT_TRACKING, max_iters=1)
    Tw1, _, i1 = oracle.opt_pose_ray_dist_sim3_numpy(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt, cfg1)
    o1, _, s1 = trk.opt_pose_ray_dist_sim3(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, v_opt, cfg1,
                                           return_stats=True)
    assert s1["iterations"] == i1["iterations"] == 1
    _check_pose(o1, Tw1, d["T_WCf"], "one iteration")
    none = np.zeros_like(v_opt)
    o2, _, s2 = trk.opt_pose_ray_dist_sim3(Xf, d["Xk"], d["T_WCf"], d["T_WCk"], Qk, none, return_stats=True)
    assert s2["iterations"] == 1 and s2["cost"] == 0.0
    assert np.abs(o2 - d["T_WCf"]).max() <= 2.4e-7
    # torch in -> torch out, same numbers
    import torch
    tw, tr = trk.opt_pose_ray_dist_sim3(torch.from_numpy(Xf).cuda(), torch.from_numpy(d["Xk"]).cuda(),
                                        d["T_WCf"], d["T_WCk"], Qk, v_opt, cfg1)
    assert tw.is_cuda and np.array_equal(tw.cpu().numpy(), o1)


def test_full_size_and_frame_tracker(mods, oracle):
    """512x384 (config 4's frame size): converges to the ground truth; FrameTracker glue."""
    trk, synth, _ = mods
    d = synth.make_track_pair(7)
    ft = trk.FrameTracker()
    res = ft.track_arrays(d["idx_f2k"], d["valid_match"], d["Xf"], d["Cf"], d["Qff"], d["Xk"], d["Ck"],
                          d["Qkf"], d["T_WCf"], d["T_WCk"])
    assert res["try_reloc"] is False and res["new_kf"] is False
    assert np.abs(res["T_WCf"] - d["T_WCf_gt"]).max() < 2e-3
    Qk, v_opt, _, (n_opt, _) = _gated(oracle, d)
    assert abs(res["match_frac"] - n_opt / len(Qk)) < 1e-12
    Tw, _, _ = oracle.opt_pose_ray_dist_sim3_numpy(d["Xf"][d["idx_f2k"]], d["Xk"], d["T_WCf"], d["T_WCk"],
                                                   Qk, v_opt)
    _check_pose(res["T_WCf"], Tw, d["T_WCf"], "full size")
    # too few matches -> relocalisation request, pose untouched (tracker.py:116-119)
    res2 = ft.track_arrays(d["idx_f2k"], np.zeros_like(d["valid_match"]), d["Xf"], d["Cf"], d["Qff"],
                           d["Xk"], d["Ck"], d["Qkf"], d["T_WCf"], d["T_WCk"])
    assert res2["try_reloc"] is True and res2["T_WCf"] is d["T_WCf"]
    # calibrated variant through the same glue
    ftc = trk.FrameTracker(dict(ft.cfg, use_calib=True))
    res3 = ftc.track_arrays(d["idx_f2k"], d["valid_match"], d["Xf"], d["Cf"], d["Qff"], d["Xk"], d["Ck"],
                            d["Qkf"], d["T_WCf"], d["T_WCk"], K=d["K4"], img_size=d["img_size"])
    # (the synthetic frame's points are not on its own pixel grid, so constraining them to pixel rays
    # moves them: the check is against the restatement of the same steps, not the ground truth)
    hw = d["img_size"]
    Xf_c = oracle.constrain_points_to_ray_numpy(hw, d["Xf"][None], d["K4"])[0]
    Xk_c = oracle.constrain_points_to_ray_numpy(hw, d["Xk"][None], d["K4"])[0]
    Tw3, _, _ = oracle.opt_pose_calib_sim3_numpy(Xf_c[d["idx_f2k"]], Xk_c, d["T_WCf"], d["T_WCk"], Qk, v_opt,
                                                 d["K4"], hw)
    _check_pose(res3["T_WCf"], Tw3, d["T_WCf"], "calib glue")
