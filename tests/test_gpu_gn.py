"""GPU parity tests for the Gauss-Newton stages (pytest -m gpu), through the C ABI.

Bar (BASELINE.json north_star): pose/scale updates within 1e-5 relative of the reference's fp64
numpy path.  Comparisons: golden vectors produced by the reference itself, the CPU oracle on seeded
graphs (incl. H, g and dx of the first iteration in fp64), and a 200-keyframe graph."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

REL = 1e-5


@pytest.fixture(scope="module")
def K():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import kernels
    assert kernels.is_available(), "CUDA extension not loaded"
    return kernels


@pytest.fixture(autouse=True, params=["f64", "mixed"])
def precision(request, K):
    """every test runs with both per-point arithmetics of the accumulate pass (include/m3s.h,
    M3S_GN_PREC_*); the 1e-5 bar on the pose update is the same for both"""
    K.set_gn_precision(request.param)
    yield request.param
    K.set_gn_precision(None)


@pytest.fixture(scope="module")
def synth():
    from mast3r_slam_b200 import synth as s
    return s


def upd_err(out, ref, Twc):
    """max |out - ref| relative to the size of the update the reference made (1e-5 bar), with one
    float32 ulp of slack for the final cast (gauss_newton.py:279)."""
    out, ref, Twc = (np.asarray(x, np.float64) for x in (out, ref, Twc))
    upd = np.abs(ref - Twc).max()
    slack = 1.2e-7 * np.abs(ref).max()
    return max(np.abs(out - ref).max() - slack, 0.0) / max(upd, 1e-30)


def gargs(g):
    return [g[k] for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q")]


@pytest.mark.parametrize("it", [1, 2, 6])
def test_gn_golden_graph(K, golden, it):
    g = golden("gn_graph")
    a = gargs(g)
    assert upd_err(K.gauss_newton_rays(*a, max_iter=it), g[f"rays_it{it}"], g["Twc"]) <= REL
    assert upd_err(K.gauss_newton_points(*a, max_iter=it), g[f"points_it{it}"], g["Twc"]) <= REL
    out = K.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K3"], g["ii"], g["jj"], g["idx"],
                               g["valid"], g["Q"], tuple(int(v) for v in g["img_size"]),
                               pixel_border=1, z_eps=0.0, sigma_pixel=1.0, sigma_depth=10.0,
                               max_iter=it)
    assert out.dtype == np.float32 and out.shape == g["Twc"].shape
    assert upd_err(out, g[f"calib_it{it}"], g["Twc"]) <= REL


def test_gn_golden_variants(K, golden):
    g = golden("gn_graph")
    a = gargs(g)
    assert upd_err(K.gauss_newton_rays(*a, max_iter=2, pin=2), g["rays_pin2"], g["Twc"]) <= REL
    out = K.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"],
                               g["valid"], g["Q"], tuple(int(v) for v in g["img_size"]),
                               pixel_border=0, z_eps=0.0, sigma_pixel=1.0, sigma_depth=0.1, max_iter=2)
    assert upd_err(out, g["calib_k4"], g["Twc"]) <= REL
    # [...,1]-shaped Cs / valid / Q (gauss_newton.py:84-89)
    out = K.gauss_newton_rays(g["Twc"], g["Xs"], g["Cs"][..., None], g["ii"], g["jj"], g["idx"],
                              g["valid"][..., None], g["Q"][..., None], max_iter=2)
    assert upd_err(out, g["rays_it2"], g["Twc"]) <= REL


def test_gn_golden_random_generator(K, golden):
    """the reference benchmark's own (incoherent, non-convergent) data: first iteration"""
    g = golden("gn_random")
    a = gargs(g)
    assert upd_err(K.gauss_newton_rays(*a, max_iter=1), g["rays_it1"], g["Twc"]) <= REL
    assert upd_err(K.gauss_newton_points(*a, max_iter=1), g["points_it1"], g["Twc"]) <= REL


def test_gn_early_outs(K, golden):
    g = golden("gn_graph")
    e = golden("gn_early")
    out = K.gauss_newton_rays(g["Twc"], g["Xs"], g["Cs"], g["ii"][:0], g["jj"][:0], g["idx"][:0],
                              g["valid"][:0], g["Q"][:0])
    assert np.array_equal(out, e["no_edges"])
    out = K.gauss_newton_rays(*gargs(g), pin=5)
    assert np.array_equal(out, e["pin_all"])
    out = K.gauss_newton_points(*gargs(g), pin=7)
    assert np.array_equal(out, g["Twc"])
    # all matches invalid: H = 1e-6 I, g = 0 -> dx = 0 -> stop, poses unchanged
    out = K.gauss_newton_rays(g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"],
                              np.zeros_like(g["valid"]), g["Q"])
    assert np.array_equal(out, g["Twc"])


@pytest.mark.parametrize("mode", ["rays", "points", "calib"])
def test_gn_first_iteration_internals(K, oracle, synth, mode, precision):
    """H, g (fp64) and dx of iteration 1 against the oracle's, through the step-wise C ABI"""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib
    g = synth.make_graph(3, 8, 3, 48, 64, num_pts=2001)  # odd N: scalar tail paths
    dev = torch.device("cuda", 0)
    plan = K.GNPlan(g["ii"], g["jj"], 8, 2001, 1, dev)
    kw = dict(sigma_pixel=1.0, sigma_depth=10.0) if mode == "calib" else {}
    if mode == "calib":
        params = K.make_params(_lib.GN_CALIB, 1, 1, 1e-4, 0.0, 1.5, 1.0, 10.0, g["K4"], g["img_size"], 0, 0.0)
        ref, dbg = oracle.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"],
                                             g["idx"], g["valid"], g["Q"], g["img_size"], max_iter=1,
                                             debug=True, **kw)
    elif mode == "rays":
        params = K.make_params(_lib.GN_RAYS, 1, 1, 1e-4, 0.0, 1.5, 0.003)
        ref, dbg = oracle.gauss_newton_rays(*gargs(g), max_iter=1, debug=True)
    else:
        params = K.make_params(_lib.GN_POINTS, 1, 1, 1e-4, 0.0, 1.5, 0.01)
        ref, dbg = oracle.gauss_newton_points(*gargs(g), max_iter=1, debug=True)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))
    _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid), p(Q),
                                     0, plan.E, None))
    _lib.check(lib.m3s_gn_assemble(plan.handle, None))
    torch.cuda.synchronize()
    H = plan.H().cpu().numpy()
    gv = plan.g().cpu().numpy()
    assert plan.F == dbg["F"] and H.shape == dbg["H"].shape
    scale = np.abs(dbg["H"]).max()
    # fp64 per-point arithmetic: round-off of a different summation order only; mixed: zero-mean
    # fp32 rounding of ~2000 terms per edge
    tol_h, tol_g, tol_dx = (1e-11, 1e-11, 1e-7) if precision == "f64" else (2e-6, 2e-6, 2e-5)
    assert np.abs(H - dbg["H"]).max() <= tol_h * scale
    assert np.abs(gv - dbg["g"]).max() <= tol_g * np.abs(dbg["g"]).max()
    _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), None))
    torch.cuda.synchronize()
    dx = plan.dx().cpu().numpy()
    assert np.abs(dx - dbg["dx"][0]).max() <= tol_dx * np.abs(dbg["dx"][0]).max()
    st = plan.status()
    assert st["applied"] == 1 and not st["failed"]
    out = torch.empty_like(Twc)
    _lib.check(lib.m3s_gn_get_poses(plan.handle, p(out), None))
    assert upd_err(out.cpu().numpy(), ref, g["Twc"]) <= REL


def test_gn_edge_sharding_is_bitwise(K, synth):
    """edge ranges accumulated separately and summed (the multi-GPU all-reduce) reproduce the
    single-pass per-edge blocks bit for bit"""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib
    g = synth.make_graph(4, 6, 2, 32, 48)
    dev = torch.device("cuda", 0)
    n = g["Xs"].shape[1]
    plan = K.GNPlan(g["ii"], g["jj"], 6, n, 1, dev)
    params = K.make_params(_lib.GN_POINTS, 1, 1, 1e-4, 0.0, 1.5, 0.01)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))

    def acc(b, e):
        _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid),
                                         p(Q), b, e, None))
        torch.cuda.synchronize()
        return plan.edge_blocks().clone()

    full = acc(0, plan.E)
    parts = acc(0, 5) + acc(5, 11) + acc(11, plan.E)
    assert torch.equal(full, parts)
    assert full.abs().sum() > 0


@pytest.mark.parametrize("cfg", [(10, 3, 2048, "rays"), (40, 12, 600, "points"), (12, 4, 1500, "calib")])
def test_gn_vs_oracle_graphs(K, oracle, synth, cfg):
    """config-3-like graphs at reduced num_pts (the numpy oracle needs minutes per iteration at full
    size); multi-panel Cholesky (dim up to 273)."""
    nkf, loops, npts, mode = cfg
    g = synth.make_graph(10 + nkf, nkf, loops, 96, 128, num_pts=npts)
    if mode == "calib":
        out = K.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"],
                                   g["valid"], g["Q"], g["img_size"], sigma_depth=10.0, max_iter=10)
        ref = oracle.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], g["ii"], g["jj"], g["idx"],
                                        g["valid"], g["Q"], g["img_size"], sigma_depth=10.0, max_iter=10)
    else:
        fn = "gauss_newton_" + mode
        out = getattr(K, fn)(*gargs(g), max_iter=10)
        ref = getattr(oracle, fn)(*gargs(g), max_iter=10)
    assert upd_err(out, ref, g["Twc"]) <= REL
    # and the solve moved the poses towards the ground truth
    assert np.abs(out - g["Twc_gt"]).max() < 0.5 * np.abs(g["Twc"] - g["Twc_gt"]).max()


def test_gn_200_keyframes(K, oracle, synth):
    """BASELINE.json config 5 topology (200 KF, 1000 directed edges, dim 1393) at reduced num_pts"""
    g = synth.make_graph(7, 200, 103, 48, 64, num_pts=512)
    assert len(g["ii"]) == 1000
    out = K.gauss_newton_points(*gargs(g), max_iter=3)
    ref = oracle.gauss_newton_points(*gargs(g), max_iter=3)
    assert upd_err(out, ref, g["Twc"]) <= REL


def test_l2_global_opt_entry_points(K, oracle, synth):
    from mast3r_slam_b200 import global_opt
    g = synth.make_graph(2, 5, 1, 32, 48)
    e = len(g["ii"]) // 2
    # split the generator's interleaved two-way rows into the FactorGraph's one-way storage
    fw, bw = slice(0, None, 2), slice(1, None, 2)
    two = global_opt.prep_two_way_edges(g["ii"][fw], g["jj"][fw], g["idx"][fw], g["idx"][bw],
                                        g["valid"][fw], g["valid"][bw], g["Q"][fw], g["Q"][bw])
    assert len(two[0]) == 2 * e and np.array_equal(two[0][e:], g["jj"][fw])
    out = global_opt.solve_GN_rays(g["Twc"], g["Xs"], g["Cs"], *two)
    ref = oracle.gauss_newton_rays(g["Twc"], g["Xs"], g["Cs"], *two, sigma_ray=0.003, max_iter=10,
                                   delta_thresh=1e-3)
    assert upd_err(out, ref, g["Twc"]) <= REL
    out = global_opt.solve_GN_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], *two, g["img_size"])
    ref = oracle.gauss_newton_calib(g["Twc"], g["Xs"], g["Cs"], g["K4"], *two, g["img_size"],
                                    sigma_pixel=1.0, sigma_depth=10.0, max_iter=10, delta_thresh=1e-3)
    assert upd_err(out, ref, g["Twc"]) <= REL


def _torch_graph_to_numpy(g):
    return dict(Twc=g["Twc"], Xs=g["Xs"].cpu().numpy(), Cs=g["Cs"].cpu().numpy(), ii=g["ii"], jj=g["jj"],
                idx=g["idx"].cpu().numpy(), valid=g["valid"].cpu().numpy().astype(bool),
                Q=g["Q"].cpu().numpy())


def test_full_size_config3_rays(K, oracle, synth):
    """BASELINE.json config 3: gauss_newton_rays, 10 keyframes / 40 directed edges, N = 196 608,
    full size against the (OpenMP) oracle"""
    import torch
    g = synth.make_graph_torch(torch.device("cuda", 0), 3, 10, 3, 384, 512)
    assert len(g["ii"]) == 40 and g["Xs"].shape[1] == 196608
    out = K.gauss_newton_rays(torch.from_numpy(g["Twc"]).cuda(), g["Xs"], g["Cs"], g["ii"], g["jj"],
                              g["idx"], g["valid"], g["Q"], max_iter=10)
    h = _torch_graph_to_numpy(g)
    ref = oracle.gauss_newton_rays(*gargs(h), max_iter=10)
    assert upd_err(out.cpu().numpy(), ref, h["Twc"]) <= REL
    assert np.abs(ref - g["Twc_gt"]).max() < 0.2 * np.abs(h["Twc"] - g["Twc_gt"]).max()


def test_full_size_config4_calib_sequence(K, oracle, synth):
    """BASELINE.json config 4: tracking-style 2-keyframe calibrated solves (keyframe pinned, frame
    free, both edge directions) at 512x384 with known intrinsics, a short synthetic sequence"""
    import torch
    for frame in range(3):
        g = synth.make_graph_torch(torch.device("cuda", 0), 40 + frame, 2, 0, 384, 512)
        assert len(g["ii"]) == 2
        out = K.gauss_newton_calib(torch.from_numpy(g["Twc"]).cuda(), g["Xs"], g["Cs"], g["K4"], g["ii"],
                                   g["jj"], g["idx"], g["valid"], g["Q"], g["img_size"], pixel_border=0,
                                   z_eps=0.0, sigma_pixel=1.0, sigma_depth=10.0, max_iter=10)
        h = _torch_graph_to_numpy(g)
        ref = oracle.gauss_newton_calib(h["Twc"], h["Xs"], h["Cs"], g["K4"], h["ii"], h["jj"], h["idx"],
                                        h["valid"], h["Q"], g["img_size"], pixel_border=0, z_eps=0.0,
                                        sigma_pixel=1.0, sigma_depth=10.0, max_iter=10)
        assert upd_err(out.cpu().numpy(), ref, h["Twc"]) <= REL
        assert np.array_equal(out.cpu().numpy()[0], h["Twc"][0])  # pinned keyframe untouched


def test_full_size_config5_points(K, oracle, synth):
    """BASELINE.json config 5 at FULL size: 200 keyframes / 1000 directed edges / 196 608 points,
    gauss_newton_points, 2 iterations against the oracle (2.4 GB of inputs)"""
    import torch
    g = synth.make_graph_torch(torch.device("cuda", 0), 0, 200, 103, 384, 512)
    assert len(g["ii"]) == 1000
    out = K.gauss_newton_points(torch.from_numpy(g["Twc"]).cuda(), g["Xs"], g["Cs"], g["ii"], g["jj"],
                                g["idx"], g["valid"], g["Q"], max_iter=2)
    h = _torch_graph_to_numpy(g)
    del g
    torch.cuda.empty_cache()
    ref = oracle.gauss_newton_points(*gargs(h), max_iter=2)
    assert upd_err(out.cpu().numpy(), ref, h["Twc"]) <= REL


@pytest.mark.parametrize("mode", ["rays", "points"])
@pytest.mark.parametrize("npts", [592, 2000, 4112])
def test_gn_streamed_partial_stages(K, oracle, synth, mode, npts):
    """N % 16 == 0 but not a multiple of the 512-point ring stage: the streamed accumulate kernel
    (cp.async.bulk ring) runs a partial last stage whose tail lanes read stale ring bytes and must
    be gated out; 592 < one stage, 2000 = 3 stages + 464, 4112 = 8 stages + 16."""
    g = synth.make_graph(21 + npts % 7, 6, 2, 96, 128, num_pts=npts)
    fn = "gauss_newton_" + mode
    out = getattr(K, fn)(*gargs(g), max_iter=6)
    ref = getattr(oracle, fn)(*gargs(g), max_iter=6)
    assert upd_err(out, ref, g["Twc"]) <= REL


@pytest.mark.parametrize("mode", ["rays", "points"])
def test_gn_gated_points_cannot_poison(K, synth, mode):
    """The packed accumulate neutralises gated-out points instead of branching around them: NaN /
    Inf coordinates, confidences and Q of points that fail the gate (valid = 0, Q <= Q_thresh,
    C <= C_thresh, idx out of range) must leave the result bit-identical, as in the reference,
    which never touches them (gauss_newton.py:129)."""
    g = synth.make_graph(5, 6, 2, 96, 128, num_pts=2048)
    fn = getattr(K, "gauss_newton_" + mode)
    clean = fn(*gargs(g), max_iter=4)
    rng = np.random.default_rng(0)
    h = {k: np.array(v, copy=True) for k, v in g.items() if isinstance(v, np.ndarray)}
    valid = h["valid"].reshape(len(h["ii"]), -1).copy()
    Q = h["Q"].reshape(valid.shape).copy()
    idx = h["idx"].reshape(valid.shape).copy()
    # per edge: points that are invalid get garbage Q / idx; a few valid ones are gated by Q instead
    bad = ~valid.astype(bool)
    Q[bad] = rng.choice([np.nan, np.inf, -np.inf, 1e30], size=int(bad.sum())).astype(np.float32)
    idx[bad] = rng.integers(-10 * 2048, 10 * 2048, size=int(bad.sum()))
    h["Q"], h["idx"] = Q.reshape(h["Q"].shape), idx.reshape(h["idx"].shape)
    # keyframe points that no valid match ever reads (as source or as target) become NaN / Inf
    n = 2048
    used = np.zeros((h["Xs"].shape[0], n), bool)
    for e, (i, j) in enumerate(zip(h["ii"], h["jj"])):
        v = valid[e].astype(bool)
        used[j, v] = True
        used[i, idx[e][v] % n] = True
    Xs = h["Xs"].reshape(-1, n, 3)
    Xs[~used] = np.nan
    h["Xs"] = Xs.reshape(h["Xs"].shape)
    out = fn(*[h[k] for k in ("Twc", "Xs", "Cs", "ii", "jj", "idx", "valid", "Q")], max_iter=4)
    assert np.array_equal(out, clean)


@pytest.mark.parametrize("num_kf,loops", [(6, 2), (24, 5), (40, 10), (200, 103)])
def test_solve_paths_agree(K, synth, monkeypatch, num_kf, loops, precision):
    """The three factorisations (launch per block step, persistent chain + workers, dataflow over
    shared-memory-resident tiles) and both back-substitutions (one CTA per 32-column block / per group of
    four) solve the same damped system: dx agrees to round-off, poses to 1e-5 of the update.  Covers one
    tile (np = 64), odd numbers of 32-column blocks (np = 288: the last tile is half padding) and the
    200-keyframe system of config 5 (44 blocks, 253 CTAs)."""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib
    if precision != "f64":
        pytest.skip("the solve is independent of the accumulate precision")
    g = synth.make_graph(11, num_kf, loops, 48, 64, num_pts=1024)
    dev = torch.device("cuda", 0)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    params = K.make_params(_lib.GN_POINTS, 1, 1, 0.0, 0.0, 1.5, 0.01)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    res = {}
    for chol, back in (("steps", "block"), ("steps", "group"), ("fused", "group"), ("tiles", "group"), ("tiles", "block"),
                       ("sparse", "group")):
        monkeypatch.setenv("M3S_GN_CHOL", chol)
        monkeypatch.setenv("M3S_GN_BACKSOLVE", back)
        plan = K.GNPlan(g["ii"], g["jj"], num_kf, 1024, 1, dev)
        _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))
        _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid), p(Q),
                                         0, plan.E, None))
        _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), None))
        torch.cuda.synchronize()
        st = plan.status()
        assert st["applied"] == 1 and not st["failed"], (chol, back, st)
        res[(chol, back)] = (plan.factor_path(), plan.dx().cpu().numpy().copy(), plan.poses().cpu().numpy().copy())
        del plan
    assert res[("steps", "block")][0] == 0 and res[("tiles", "group")][0] == 2 and res[("sparse", "group")][0] == 3
    ref_dx, ref_pose = res[("steps", "block")][1:]
    scale = np.abs(ref_dx).max()
    for key, (_, dx, pose) in res.items():
        assert np.abs(dx - ref_dx).max() <= 1e-8 * scale, key
        assert np.abs(pose - ref_pose).max() <= 1e-9 * max(np.abs(ref_pose).max(), 1.0), key
    # the grouped back-substitution reads the same factor: bitwise equal to the per-block one
    assert np.array_equal(res[("tiles", "group")][1], res[("tiles", "block")][1]) or \
        np.abs(res[("tiles", "group")][1] - res[("tiles", "block")][1]).max() <= 1e-12 * scale


def test_sparse_front_selection(K, synth, monkeypatch, precision):
    """The sparse front (levels of exactly eliminated keyframes + dense tail) is chosen from the graph alone:
    taken for the window-2 chain + loop closures of config 5, not for a complete graph (nothing to eliminate
    cheaply) nor for systems of one tile; M3S_GN_SPARSE=0 keeps everything dense."""
    import torch
    if precision != "f64":
        pytest.skip("plan-level property")
    dev = torch.device("cuda", 0)
    monkeypatch.delenv("M3S_GN_CHOL", raising=False)
    monkeypatch.delenv("M3S_GN_SPARSE", raising=False)
    rng = np.random.default_rng(5)
    ii, jj = synth.graph_edges(200, 103, rng)
    plan = K.GNPlan(ii, jj, 200, 1024, 1, dev)
    info = plan.sparse_info()
    assert info["levels"] >= 3 and info["eliminated"] + info["tail"] == plan.dim // 7
    assert info["tail_np"] <= plan.ld // 2 and plan.factor_path() == 3
    del plan
    n = 30
    a, b = np.triu_indices(n, 1)
    ci, cj = np.concatenate([a, b]).astype(np.int32), np.concatenate([b, a]).astype(np.int32)
    plan = K.GNPlan(ci, cj, n, 1024, 1, dev)
    assert plan.sparse_info()["levels"] == 0 and plan.factor_path() == 2
    del plan
    ii, jj = synth.graph_edges(8, 1, rng)
    plan = K.GNPlan(ii, jj, 8, 1024, 1, dev)
    assert plan.sparse_info()["levels"] == 0
    del plan
    monkeypatch.setenv("M3S_GN_SPARSE", "0")
    ii, jj = synth.graph_edges(200, 103, rng)
    plan = K.GNPlan(ii, jj, 200, 1024, 1, dev)
    assert plan.sparse_info()["levels"] == 0 and plan.factor_path() == 2


@pytest.mark.parametrize("num_kf,loops", [(40, 10), (120, 40)])
def test_sparse_front_iterations_and_stop(K, synth, monkeypatch, num_kf, loops):
    """Whole solves through the sparse front against the dense dataflow path: several iterations (poses to
    1e-9 of the pose scale), the same iteration count when the ||dx|| test stops the loop, and an untouched
    pose array when the first step is already below the threshold (stop BEFORE applying, gauss_newton.py:263)."""
    g = synth.make_graph(21, num_kf, loops, 48, 64, num_pts=2048)
    a = gargs(g)
    out = {}
    for chol in ("tiles", "sparse"):
        monkeypatch.setenv("M3S_GN_CHOL", chol)
        K._plan_cache.clear()
        out[chol] = (K.gauss_newton_points(*a, max_iter=6, delta_thresh=0.0),
                     K.gauss_newton_points(*a, max_iter=10, delta_thresh=1e-4),
                     K.gauss_newton_points(*a, max_iter=10, delta_thresh=1e9))
        plan = K.get_plan(g["ii"], g["jj"], num_kf, 2048, 1, __import__("torch").device("cuda", 0))
        assert plan.factor_path() == (3 if chol == "sparse" else 2)
        out[chol] += (plan.status()["attempted"],)
    K._plan_cache.clear()
    scale = np.abs(g["Twc"]).max()
    for k in range(2):
        assert np.abs(out["sparse"][k].astype(np.float64) - out["tiles"][k]).max() <= 1e-6 * scale
        assert upd_err(out["sparse"][k], out["tiles"][k], g["Twc"]) <= 1e-6
    assert np.array_equal(out["sparse"][2], g["Twc"].astype(np.float32)) and np.array_equal(out["tiles"][2], out["sparse"][2])
    assert out["sparse"][3] == out["tiles"][3]


def test_sparse_front_long_trajectory(K, synth, monkeypatch, precision):
    """500 keyframes / 30 loop closures: the dense system (np = 3520, 55 tile rows) does not fit the co-resident
    tile grid and would take the launch-per-step path; the sparse front leaves a tail that does.  Same dx as the
    launch-per-step factorisation of the whole system."""
    import ctypes as C
    import torch
    from mast3r_slam_b200 import _lib
    if precision != "f64":
        pytest.skip("the solve is independent of the accumulate precision")
    g = synth.make_graph(31, 500, 30, 24, 32, num_pts=256)
    dev = torch.device("cuda", 0)
    t = lambda a, dt: torch.from_numpy(np.ascontiguousarray(a)).to(dev, dt)  # noqa: E731
    Twc, Xs, Cs = t(g["Twc"], torch.float32), t(g["Xs"], torch.float32), t(g["Cs"], torch.float32)
    idx, valid, Q = t(g["idx"], torch.int32), t(g["valid"].astype(np.uint8), torch.uint8), t(g["Q"], torch.float32)
    params = K.make_params(_lib.GN_POINTS, 1, 1, 0.0, 0.0, 1.5, 0.01)
    lib = _lib.load()
    p = lambda x: C.c_void_p(x.data_ptr())  # noqa: E731
    res = {}
    for chol in ("steps", None):
        if chol:
            monkeypatch.setenv("M3S_GN_CHOL", chol)
        else:
            monkeypatch.delenv("M3S_GN_CHOL", raising=False)
        plan = K.GNPlan(g["ii"], g["jj"], 500, 256, 1, dev)
        _lib.check(lib.m3s_gn_set_poses(plan.handle, p(Twc), None))
        _lib.check(lib.m3s_gn_accumulate(plan.handle, C.byref(params), p(Xs), p(Cs), p(idx), p(valid), p(Q),
                                         0, plan.E, None))
        _lib.check(lib.m3s_gn_solve_update(plan.handle, C.byref(params), None))
        torch.cuda.synchronize()
        st = plan.status()
        assert st["applied"] == 1 and not st["failed"], (chol, st)
        res[chol] = (plan.factor_path(), plan.sparse_info(), plan.dx().cpu().numpy().copy())
        del plan
    assert res["steps"][0] == 0 and res[None][0] == 3, (res["steps"][:2], res[None][:2])
    assert res[None][1]["tail"] < 150
    scale = np.abs(res["steps"][2]).max()
    assert np.abs(res[None][2] - res["steps"][2]).max() <= 1e-8 * scale
