"""Golden vectors for the rows of SURVEY.md s8 whose reference code needs MLX (a1, a4, f1-f3):
the UNMODIFIED reference modules are imported from /root/reference/src and executed on top of the
numpy-backed ``mlx.core`` stand-in in tests/golden/mlx_shim (the container has no MLX / Metal).

    python tests/golden/gen_golden_mlx.py

Functions executed (reference file:line):
  * matching.prep_for_iter_proj (matching.py:134-175) + image.img_gradient (image.py:9-34)
  * matching.match_iterative_proj, whole call (matching.py:339-461: prep, kernels.iter_proj ->
    numpy fallback, kernels.refine_matches -> numpy fallback, tail)
  * matching.match / match_simple, the default route (matching.py:12-90)
  * geometry.constrain_points_to_ray (geometry.py:273-302)
  * global_opt.FactorGraph.add_factors / _prep_two_way_edges / solve_GN_calib on real
    frame.Keyframes / frame.Frame objects with a stub matcher (global_opt.py:49-154, 212-268, 491-545)
  * tracker.FrameTracker._get_points_poses / _opt_pose_ray_dist_sim3 / _opt_pose_calib_sim3 and the
    whole track() with a stub matcher (tracker.py:51-406)
The stand-in computes in float32 like MLX does; what it cannot reproduce is MLX's internal order of
multi-term reductions (numpy's is used).  Fixtures are small (.npz) and committed.
"""
import copy
import os
import sys
import warnings

import numpy as np

sys.dont_write_bytecode = True  # /root/reference is read-only
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, "/root/reference/src")
sys.path.insert(0, os.path.join(HERE, "mlx_shim"))
sys.path.insert(0, ROOT)
warnings.filterwarnings("ignore")

import mlx.core as mx  # noqa: E402  (the stand-in)
from mlx_mast3r_slam import config as rcfg, frame as rframe, geometry as rgeo  # noqa: E402
from mlx_mast3r_slam import global_opt as rgo, matching as rmatch, tracker as rtrack  # noqa: E402
from mlx_mast3r_slam.liegroups import Sim3  # noqa: E402

import mast3r_slam_b200.synth as synth  # noqa: E402

assert mx.__file__.startswith(HERE), "the numpy stand-in must shadow any real mlx"


def save(name, **kw):
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **{k: np.asarray(v) for k, v in kw.items()})
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def A(x, dt=None):
    return mx.array(np.asarray(x), dtype=dt)


def N(x):
    return np.array(np.asarray(x))


def set_ref_config(**sections):
    cfg = copy.deepcopy(rcfg.DEFAULT_CONFIG)
    for k, v in sections.items():
        cfg.setdefault(k, {}).update(v)
    rcfg.config = cfg  # what load_config() assigns (config.py:41); get_config() returns it
    return cfg


def gen_prep_and_match():
    h, w = 24, 32
    pair = synth.make_pair(4, h, w, 24, "dyadic")
    X11, X21, D11, D21 = (pair[k][None] for k in ("X11", "X21", "D11", "D21"))
    rng = np.random.default_rng(5)
    # initial correspondences: identity jittered by a few pixels, clipped into the image
    uv = np.stack(np.meshgrid(np.arange(w), np.arange(h)), -1).reshape(-1, 2)
    j = uv + rng.integers(-2, 3, uv.shape)
    idx_init = (np.clip(j[:, 1], 0, h - 1) * w + np.clip(j[:, 0], 0, w - 1)).astype(np.int32)[None]
    out = {}
    for tag, init in (("none", None), ("init", idx_init)):
        r, p, p0 = rmatch.prep_for_iter_proj(A(X11), A(X21), None if init is None else A(init))
        out.update({f"rays_{tag}": N(r), f"pts_{tag}": N(p), f"p_init_{tag}": N(p0)})
    gx, gy = rmatch.img_gradient(mx.transpose(A(X11), (0, 3, 1, 2)))
    out["img_gradient_gx"], out["img_gradient_gy"] = N(gx), N(gy)
    # whole match_iterative_proj under three configurations
    set_ref_config(matching={"use_simple": False})
    for tag, init in (("none", None), ("init", idx_init)):
        idx, valid = rmatch.match_iterative_proj(A(X11), A(X21), A(D11), A(D21), None if init is None else A(init))
        out[f"mip_idx_{tag}"], out[f"mip_valid_{tag}"] = N(idx), N(valid)
    idx, valid = rmatch.match(A(X11), A(X21), A(D11), A(D21))
    assert np.array_equal(N(idx), out["mip_idx_none"])
    set_ref_config(matching={"use_simple": False, "refine_radius": 2, "dist_thresh": 0.3, "max_iter": 4})
    idx, valid = rmatch.match_iterative_proj(A(X11), A(X21), A(D11), A(D21))
    out["mip_idx_r2"], out["mip_valid_r2"] = N(idx), N(valid)
    set_ref_config(matching={"use_simple": False, "use_refine": False})
    idx, valid = rmatch.match_iterative_proj(A(X11), A(X21), A(D11), A(D21))
    out["mip_idx_norefine"], out["mip_valid_norefine"] = N(idx), N(valid)
    # the default route (config.py:65 use_simple=True -> match_simple)
    set_ref_config()
    assert rcfg.get_config()["matching"]["use_simple"] is True
    for tag, init in (("none", None), ("init", idx_init)):
        idx, valid = rmatch.match(A(X11), A(X21), A(D11), A(D21), None if init is None else A(init))
        out[f"simple_idx_{tag}"], out[f"simple_valid_{tag}"] = N(idx), N(valid)
    set_ref_config(matching={"dist_thresh": 0.25})
    idx, valid = rmatch.match(A(X11), A(X21), A(D11), A(D21), A(idx_init))
    out["simple_valid_thr025"] = N(valid)
    set_ref_config()
    save("mlx_match", X11=X11, X21=X21, D11=D11, D21=D21, idx_init=idx_init, **out)


def _keyframes(g, h, w, n_updates=None):
    """real reference Keyframes / Frame objects filled from a synthetic graph"""
    kfs = rframe.Keyframes(h, w)
    k = g["Xs"].shape[0]
    for i in range(k):
        f = rframe.Frame(frame_id=i, img=mx.zeros((3, h, w)), img_shape=A([[h, w]]), img_true_shape=A([[h, w]]),
                         T_WC=Sim3(A(g["Twc"][i][None])))
        nu = 1 if n_updates is None else int(n_updates[i])
        f.X_canon = A(g["Xs"][i])
        f.C = A(g["Cs"][i][:, None] * nu)
        f.N = nu
        f.N_updates = nu
        kfs.append(f)
    return kfs


def gen_graph():
    h, w = 20, 28  # non-square on purpose: solve_GN_calib passes (H, W) where (W, H) is unpacked
    n = h * w
    g = synth.make_graph(2, 5, 1, h, w)
    n_updates = np.array([1, 2, 1, 3, 1])
    K3 = np.array([[g["K4"][0], 0, g["K4"][2]], [0, g["K4"][1], g["K4"][3]], [0, 0, 1]], np.float32)
    out = {}
    # constrain_points_to_ray, both K forms the reference accepts
    out["ray_Xs"] = N(rgeo.constrain_points_to_ray((h, w), A(g["Xs"]), A(K3)))
    # add_factors: undirected pairs from the synthetic graph, matcher outputs synthesised
    E = len(g["ii"]) // 2
    ii_u, jj_u = g["ii"][:E], g["jj"][:E]
    rng = np.random.default_rng(9)
    Qii, Qjj, Qji, Qij = (rng.uniform(0.8, 4.0, (E, n, 1)).astype(np.float32) for _ in range(4))
    idx_i2j, idx_j2i = g["idx"][:E], g["idx"][E:]
    vj, vi = g["valid"][:E, :, None].copy(), g["valid"][E:, :, None].copy()
    # one non-consecutive pair with almost no matches -> filtered by min_match_frac
    nc = [e for e in range(E) if ii_u[e] != jj_u[e] - 1]
    assert nc, "need a loop-closure pair"
    vj[nc[0]] = False
    vj[nc[0], :3] = True
    set_ref_config(local_opt={"max_iters": 4, "sigma_depth": 10.0})
    kfs = _keyframes(g, h, w, n_updates)
    fg = rgo.FactorGraph(None, kfs, K=A(K3))
    half = E // 2

    def matcher(lo, hi):
        def fn(model, feat_i, pos_i, feat_j, pos_j, shape_i, shape_j):
            s = slice(lo, hi)
            return (A(idx_i2j[s]), A(idx_j2i[s]), A(vj[s]), A(vi[s]), A(Qii[s]), A(Qjj[s]), A(Qji[s]), A(Qij[s]))
        return fn

    for f in kfs._frames:  # add_factors concatenates feat / pos before calling the matcher
        f.feat, f.pos = mx.zeros((1, 1, 1)), mx.zeros((1, 1, 2))
    r1 = fg.add_factors(list(ii_u[:half]), list(jj_u[:half]), 0.1, matcher(0, half))
    r2 = fg.add_factors(list(ii_u[half:]), list(jj_u[half:]), 0.1, matcher(half, E))
    # a relocalisation attempt that contains the bad pair is refused as a whole
    r3 = fg.add_factors([int(ii_u[nc[0]])], [int(jj_u[nc[0]])], 0.1, matcher(nc[0], nc[0] + 1), is_reloc=True)
    out.update(add_ret=np.array([r1, r2, r3]), fg_ii=N(fg.ii), fg_jj=N(fg.jj), fg_idx_ii2jj=N(fg.idx_ii2jj),
               fg_idx_jj2ii=N(fg.idx_jj2ii), fg_valid_match_j=N(fg.valid_match_j),
               fg_valid_match_i=N(fg.valid_match_i), fg_Q_ii2jj=N(fg.Q_ii2jj), fg_Q_jj2ii=N(fg.Q_jj2ii))
    tw = fg._prep_two_way_edges()
    for k, v in zip(("ii", "jj", "idx", "valid", "Q"), tw):
        out["two_way_" + k] = N(v)
    out["unique_kf"] = N(fg.get_unique_kf_idx())
    fg.solve_GN_calib()
    out["poses_after_calib"] = np.stack([N(f.T_WC.data)[0] for f in kfs._frames])
    set_ref_config()
    save("mlx_graph", Twc=g["Twc"], Xs=g["Xs"], Cs=g["Cs"], n_updates=n_updates, K3=K3, hw=np.array([h, w]),
         ii_u=ii_u, jj_u=jj_u, idx_i2j=idx_i2j, idx_j2i=idx_j2i, valid_j=vj, valid_i=vi, Qii=Qii, Qjj=Qjj,
         Qji=Qji, Qij=Qij, min_match_frac=np.float64(0.1), **out)


def gen_track():
    out = {}
    for tag, (seed, h, w) in (("a", (0, 24, 32)), ("b", (3, 32, 40))):
        d = synth.make_track_pair(seed, h, w)
        n = h * w
        K3 = np.array([[d["K4"][0], 0, d["K4"][2]], [0, d["K4"][1], d["K4"][3]], [0, 0, 1]], np.float32)
        for use_calib in (False, True):
            t = f"{tag}_{'calib' if use_calib else 'ray'}"
            cfg = set_ref_config(tracking={"use_calib": use_calib})
            kfs = rframe.Keyframes(h, w)
            mk = lambda i, X, Cc, T: rframe.Frame(  # noqa: E731
                frame_id=i, img=mx.zeros((3, h, w)), img_shape=A([[h, w]]), img_true_shape=A([[h, w]]),
                T_WC=Sim3(A(T[None])), X_canon=A(X), C=A(Cc[:, None]), N=1, N_updates=1, K=A(K3))
            kf = mk(0, d["Xk"], d["Ck"], d["T_WCk"])
            fr = mk(1, d["Xf"], d["Cf"], d["T_WCf"])
            fr.N = 0  # track() calls update_pointmap(Xff, Cff) first: N == 0 -> adopts them
            fr.N_updates = 0
            kfs.append(kf)
            trk = rtrack.FrameTracker(None, kfs)

            def match_fn(model, frame, keyframe, idx_i2j_init=None):
                return (A(d["idx_f2k"][None]), A(d["valid_match"].reshape(1, n, 1)), A(d["Xf"]), A(d["Cf"][:, None]),
                        A(d["Qff"][None, :, None]), A(d["Xk"]), A(d["Ck"][:, None]), A(d["Qkf"][None, :, None]))

            # the solver entry points on their own (inputs exactly as track() prepares them)
            fr2 = copy.copy(fr)
            fr2.update_pointmap(A(d["Xf"]), A(d["Cf"][:, None]))
            img_size = (h, w)
            Xf, Xk, T_WCf, T_WCk, Cf, Ck, meas_k, valid_meas_k = trk._get_points_poses(
                fr2, kf, A(d["idx_f2k"]), img_size, use_calib, A(K3) if use_calib else None)
            Qk = mx.sqrt(A(d["Qff"])[A(d["idx_f2k"])] * A(d["Qkf"]))[:, None]
            tc = cfg["tracking"]
            valid_opt = A(d["valid_match"].reshape(n, 1)) & (Cf > tc["C_conf"]) & (Ck > tc["C_conf"]) & (Qk > tc["Q_conf"])
            if use_calib:
                T_new, T_rel = trk._opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid_opt, meas_k, valid_meas_k,
                                                        A(K3), img_size)
                out[f"{t}_meas_k"], out[f"{t}_valid_meas_k"] = N(meas_k), N(valid_meas_k)
            else:
                T_new, T_rel = trk._opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid_opt)
            out.update({f"{t}_Xf_in": N(Xf), f"{t}_Xk_in": N(Xk), f"{t}_Qk": N(Qk), f"{t}_valid_opt": N(valid_opt),
                        f"{t}_T_WCf": N(T_new.data)[0], f"{t}_T_CkCf": N(T_rel.data)[0]})
            # and the whole track()
            new_kf, match_info, try_reloc = trk.track(fr, match_fn)
            out.update({f"{t}_track_new_kf": np.array(new_kf), f"{t}_track_try_reloc": np.array(try_reloc),
                        f"{t}_track_T_WCf": N(fr.T_WC.data)[0]})
        out.update({f"{tag}_{k}": d[k] for k in ("Xf", "Xk", "Cf", "Ck", "Qff", "Qkf", "idx_f2k", "valid_match",
                                                 "T_WCf", "T_WCk", "K4")})
        out[f"{tag}_hw"] = np.array([h, w])
    set_ref_config()
    save("mlx_track", **out)


if __name__ == "__main__":
    gen_prep_and_match()
    gen_graph()
    gen_track()
