"""GPU parity tests for the factor-graph host steps: constrain_points_to_ray,
the add_factors confidence / match-fraction step, FactorGraph on device-resident storage, and
symmetric matching.  Everything goes through the C ABI; the checker is the numpy restatement of the
reference lines in oracle/oracle.py (the reference code itself is MLX and cannot be imported)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def mods():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import kernels, global_opt, matching, synth, config
    assert kernels.is_available()
    return kernels, global_opt, matching, synth, config


def test_constrain_points_to_ray(mods, oracle):
    K_, _, _, synth, _ = mods
    h, w = 37, 52
    rng = np.random.default_rng(0)
    Xs = rng.normal(size=(3, h * w, 3)).astype(np.float32)
    Xs[..., 2] = np.abs(Xs[..., 2]) + 0.5
    Xs[0, 5, 2] = 0.0
    Xs[1, 7, 2] = -1.0
    for K in (np.array([420.0, 415.0, 25.5, 18.25], np.float32),
              np.array([[100.0, 0, 26.0], [0, 101.0, 18.0], [0, 0, 1]], np.float32)):
        out = K_.constrain_points_to_ray((h, w), Xs, K)
        ref = oracle.constrain_points_to_ray_numpy((h, w), Xs, K)
        assert out.dtype == np.float32 and np.array_equal(out, ref)
    import torch
    t = K_.constrain_points_to_ray((h, w), torch.from_numpy(Xs).cuda(), K)
    assert t.is_cuda and np.array_equal(t.cpu().numpy(), ref)


def _match_out(rng, e, n, frac_valid):
    idx_i2j = rng.integers(0, n, (e, n)).astype(np.int32)
    idx_j2i = rng.integers(0, n, (e, n)).astype(np.int32)
    vmj = rng.random((e, n, 1)) < np.asarray(frac_valid).reshape(e, 1, 1)
    vmi = rng.random((e, n, 1)) < 0.9
    Q = [rng.uniform(0.5, 4.0, (e, n, 1)).astype(np.float32) for _ in range(4)]
    return (idx_i2j, idx_j2i, vmj, vmi, *Q)


def test_edge_confidence(mods, oracle):
    K_ = mods[0]
    rng = np.random.default_rng(1)
    e, n = 5, 3000
    mo = _match_out(rng, e, n, [0.9] * e)
    Qj, valid, frac = K_.edge_confidence(mo[4], mo[6], mo[0], mo[2], 1.5)
    bi = np.broadcast_to(np.arange(e)[:, None], (e, n))
    Qj_ref = np.sqrt(mo[4][bi, mo[0]] * mo[6])
    assert np.array_equal(Qj, Qj_ref)
    v_ref = mo[2] & (Qj_ref > 1.5)
    assert np.array_equal(valid, v_ref)
    assert np.array_equal(frac, v_ref.astype(np.float32).sum((1, 2)) / n)
    # out-of-range correspondence (an IndexError in the reference): Qj = 0, invalid
    bad = mo[0].copy(); bad[0, 0] = n; bad[1, 1] = -1
    Qj2, v2, _ = K_.edge_confidence(mo[4], mo[6], bad, mo[2], 1.5)
    assert Qj2[0, 0, 0] == 0 and not v2[0, 0, 0] and Qj2[1, 1, 0] == 0 and not v2[1, 1, 0]
    # empty
    Qj0, v0, f0 = K_.edge_confidence(mo[4][:0], mo[6][:0], mo[0][:0], mo[2][:0], 1.5)
    assert Qj0.shape == (0, n, 1) and f0.shape == (0,)


@pytest.mark.parametrize("is_reloc", [False, True])
def test_factor_graph_add_factors(mods, oracle, is_reloc):
    import torch
    K_, G, _, synth, config = mods
    rng = np.random.default_rng(2)
    n = 40 * 56
    store = G.KeyframeStore(n, capacity=2)
    for k in range(7):
        store.append(rng.normal(size=(n, 3)), rng.uniform(1, 9, (n, 1)), np.r_[rng.normal(size=3), 0, 0, 0, 1, 1.0], 3)
    assert len(store) == 7 and store.X.shape[0] >= 7
    cfg = dict(config.get_config()["local_opt"])
    fg = G.FactorGraph(store, cfg=cfg, capacity=1)
    # batch 1: consecutive edges with a poor match fraction are kept; non-consecutive ones are dropped
    ii, jj = [0, 1, 0, 2], [1, 2, 3, 5]
    mo = _match_out(rng, 4, n, [0.9, 0.01, 0.02, 0.9])
    ref = oracle.add_factors_numpy(ii, jj, mo, cfg["Q_conf"], 0.3, is_reloc)
    ok = fg.add_factors(ii, jj, 0.3, mo, is_reloc)
    assert ok == (ref is not None)
    if ref is None:
        assert len(fg) == 0
        return
    names = ("ii", "jj", "idx_ii2jj", "idx_jj2ii", "valid_match_j", "valid_match_i", "Q_ii2jj", "Q_jj2ii")
    for name, r in zip(names, ref):
        got = getattr(fg, name).cpu().numpy()
        assert np.array_equal(got.reshape(r.shape), r), name
    # batch 2 through a callable, storage growth, two-way views
    mo2 = _match_out(rng, 2, n, [0.8, 0.7])
    assert fg.add_factors([3, 4], [4, 6], 0.3, lambda a, b: mo2)
    ref2 = oracle.add_factors_numpy([3, 4], [4, 6], mo2, cfg["Q_conf"], 0.3)
    assert len(fg) == len(ref[0]) + len(ref2[0])
    ii2, jj2, idx2, valid2, Q2 = fg._prep_two_way_edges()
    e = len(fg)
    assert ii2.shape == (2 * e,) and idx2.shape == (2 * e, n)
    assert torch.equal(ii2[0::2], jj2[1::2]) and torch.equal(jj2[0::2], ii2[1::2])
    assert np.array_equal(idx2[0::2].cpu().numpy(), np.concatenate([ref[2], ref2[2]]))
    assert np.array_equal(Q2[1::2].cpu().numpy()[..., None], np.concatenate([ref[7], ref2[7]]))
    assert np.array_equal(fg.get_unique_kf_idx().cpu().numpy(), np.unique(np.r_[ref[0], ref[1], ref2[0], ref2[1]]))


def test_factor_graph_solves(mods, oracle):
    """FactorGraph.solve_GN_rays / solve_GN_calib on device-resident storage == the oracle run on
    the reference's [all i->j | all j->i] concatenation (global_opt.py:147-154), pinned poses
    untouched, free poses written back."""
    import torch
    K_, G, _, synth, config = mods
    h, w = 48, 64
    g = synth.make_graph(3, 6, 2, h, w)
    n = h * w
    ii_d, jj_d = g["ii"], g["jj"]
    # the synthetic graph lists every pair both ways: recover the undirected pairs i < j
    fwd = [k for k in range(len(ii_d)) if ii_d[k] < jj_d[k]]
    rev = {(int(ii_d[k]), int(jj_d[k])): k for k in range(len(ii_d))}
    bwd = [rev[(int(jj_d[k]), int(ii_d[k]))] for k in fwd]
    cfg = dict(config.get_config()["local_opt"]); cfg["max_iters"] = 4
    for mode in ("rays", "calib"):
        store = G.KeyframeStore(n, capacity=4)
        for k in range(g["Twc"].shape[0]):
            store.append(g["Xs"][k], g["Cs"][k] * 2.0, g["Twc"][k], 2)
        fg = G.FactorGraph(store, K=g["K4"], img_size=(h, w), cfg=cfg, calib_img_size_order="wh")
        ones = np.ones((len(fwd), n, 1), np.float32)
        Qf, Qb = g["Q"][fwd].reshape(-1, n, 1), g["Q"][bwd].reshape(-1, n, 1)
        # Qj = sqrt(Qii[idx] * Qji): feed Qii = 1 and Qji = Q^2 so the stored confidence equals Q
        mo = (g["idx"][fwd], g["idx"][bwd], g["valid"][fwd].reshape(-1, n, 1), g["valid"][bwd].reshape(-1, n, 1),
              ones, ones, Qf * Qf, Qb * Qb)
        assert fg.add_factors(ii_d[fwd].tolist(), jj_d[fwd].tolist(), 0.0, mo)
        Qs = np.concatenate([np.sqrt(Qf * Qf), np.sqrt(Qb * Qb)])[..., 0]
        cat = lambda a: np.concatenate([a[fwd], a[bwd]])
        args = (g["Twc"], None, g["Cs"], cat(ii_d), cat(jj_d), cat(g["idx"]), cat(g["valid"]), Qs)
        if mode == "rays":
            fg.solve_GN_rays()
            ref = oracle.gauss_newton_rays(args[0], g["Xs"], *args[2:], sigma_ray=cfg["sigma_ray"],
                                           C_thresh=cfg["C_conf"], Q_thresh=cfg["Q_conf"], max_iter=4,
                                           delta_thresh=cfg["delta_norm"], pin=cfg["pin"])
        else:
            fg.solve_GN_calib()
            Xc = oracle.constrain_points_to_ray_numpy((h, w), g["Xs"], g["K4"])
            ref = oracle.gauss_newton_calib(args[0], Xc, args[2], g["K4"], *args[3:], (w, h),
                                            sigma_pixel=cfg["sigma_pixel"], sigma_depth=cfg["sigma_depth"],
                                            C_thresh=cfg["C_conf"], Q_thresh=cfg["Q_conf"], max_iter=4,
                                            delta_thresh=cfg["delta_norm"], pin=cfg["pin"])
        out = store.T_WC[:len(store)].cpu().numpy()
        upd = np.abs(ref - g["Twc"]).max()
        assert upd > 1e-4
        assert np.abs(out.astype(np.float64) - ref).max() <= 1e-5 * upd + 2e-7, mode
        assert np.array_equal(out[:cfg["pin"]], g["Twc"][:cfg["pin"]])


def test_match_symmetric(mods, oracle):
    K_, _, M, synth, config = mods
    config.set_config({"matching": {"use_simple": False}})
    a, b = synth.make_pair(4, 48, 64), synth.make_pair(5, 48, 64)
    st = lambda k: np.stack([a[k], b[k]])
    # direction i->j uses (X11, X21); the reverse direction swaps the roles of the two views
    i2j, j2i, vj, vi = M.match_symmetric(st("X11"), st("X21"), st("X21"), st("X11"),
                                         st("D11"), st("D21"), st("D21"), st("D11"))
    r_i2j, r_vj = oracle.match_iterative_proj(st("X11"), st("X21"), st("D11"), st("D21"))
    r_j2i, r_vi = oracle.match_iterative_proj(st("X21"), st("X11"), st("D21"), st("D11"))
    assert np.array_equal(i2j, r_i2j) and np.array_equal(vj, r_vj)
    assert np.array_equal(j2i, r_j2i) and np.array_equal(vi, r_vi)
