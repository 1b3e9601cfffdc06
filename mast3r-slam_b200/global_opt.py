"""L2 Gauss-Newton host: the array-level part of the reference's ``global_opt.FactorGraph``.

Only what sits on the hot path is restated: the two-way edge
concatenation (global_opt.py:147-154) and the solver entry points (solve_GN_rays :168-210,
solve_GN_calib :212-268, _gauss_newton_calib :491-545) as functions over arrays.  Keyframe
bookkeeping (add_factors, Keyframes, retrieval) is SLAM logic and out of scope.  Unlike the
reference (global_opt.py:543-545) solver errors are raised, not swallowed.
"""
from __future__ import annotations

import numpy as np
import torch

from . import kernels
from .config import get_config


def _cat(a, b):
    if isinstance(a, torch.Tensor):
        return torch.cat([a, b])
    return np.concatenate([np.asarray(a), np.asarray(b)])


def prep_two_way_edges(ii, jj, idx_ii2jj, idx_jj2ii, valid_match_j, valid_match_i, Q_ii2jj, Q_jj2ii):
    """global_opt.py:147-154: directed rows (i->j) followed by (j->i)."""
    return (_cat(ii, jj), _cat(jj, ii), _cat(idx_ii2jj, idx_jj2ii),
            _cat(valid_match_j, valid_match_i), _cat(Q_ii2jj, Q_jj2ii))


def solve_GN_rays(T_WCs, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, cfg=None):
    """global_opt.py:168-210 with a working inner solver (the reference's MLX inner loop is a
    stub): poses [K,8] -> updated poses [K,8] via kernels.gauss_newton_rays."""
    cfg = cfg or get_config()["local_opt"]
    return kernels.gauss_newton_rays(
        T_WCs, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, cfg["sigma_ray"], cfg["sigma_dist"],
        cfg["C_conf"], cfg["Q_conf"], cfg["max_iters"], cfg["delta_norm"], cfg["pin"])


def solve_GN_calib(T_WCs, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, img_size, cfg=None):
    """global_opt.py:212-268 / :491-545.  `Xs` must already be constrained to the pixel rays
    (geometry.constrain_points_to_ray is caller-side prep)."""
    cfg = cfg or get_config()["local_opt"]
    return kernels.gauss_newton_calib(
        T_WCs, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q_ii2jj, img_size,
        cfg.get("pixel_border", 0), cfg.get("depth_eps", 0.0), cfg["sigma_pixel"],
        cfg["sigma_depth"], cfg["C_conf"], cfg["Q_conf"], cfg["max_iters"], cfg["delta_norm"],
        cfg["pin"])


# ---------------------------------------------------------------------------------------------
# FactorGraph on device-resident storage
# ---------------------------------------------------------------------------------------------
class KeyframeStore:
    """What FactorGraph reads from the reference's ``Keyframes`` (frame.py:146-260), as preallocated
    CUDA tensors: canonical pointmap ``X [K,N,3]``, average confidence ``Cavg [K,N]`` (= C / N_updates,
    frame.py:135-143), pose ``T_WC [K,8]`` (t, q_xyzw, s).  Rows are appended in place; capacity
    doubles when exhausted, so solves read the tensors without re-stacking (the reference rebuilds
    Xs / Cs / T_WCs with mx.stack on every solve, global_opt.py:156-166)."""

    def __init__(self, n_points: int, device=None, capacity: int = 16):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        self.n_points = int(n_points)
        self._k = 0
        self._alloc(capacity)

    def _alloc(self, cap):
        old = getattr(self, "X", None)
        X = torch.empty((cap, self.n_points, 3), dtype=torch.float32, device=self.device)
        Cv = torch.empty((cap, self.n_points), dtype=torch.float32, device=self.device)
        T = torch.empty((cap, 8), dtype=torch.float32, device=self.device)
        if old is not None:
            X[:self._k] = self.X[:self._k]
            Cv[:self._k] = self.Cavg[:self._k]
            T[:self._k] = self.T_WC[:self._k]
        self.X, self.Cavg, self.T_WC, self._cap = X, Cv, T, cap

    def __len__(self):
        return self._k

    def append(self, X_canon, C, T_WC, n_updates: int = 1) -> int:
        """Add a keyframe; C is the accumulated confidence, n_updates its update count."""
        if self._k == self._cap:
            self._alloc(2 * self._cap)
        k = self._k
        self.set(k, X_canon, C, T_WC, n_updates)
        self._k += 1
        return k

    def set(self, k, X_canon, C, T_WC=None, n_updates: int = 1):
        as_t = lambda a: torch.as_tensor(a, dtype=torch.float32).to(self.device)
        self.X[k] = as_t(X_canon).reshape(self.n_points, 3)
        self.Cavg[k] = as_t(C).reshape(self.n_points) / float(n_updates)
        if T_WC is not None:
            self.T_WC[k] = as_t(T_WC).reshape(8)

    def update_T_WCs(self, poses, idx):
        """frame.py Keyframes.update_T_WCs: write back optimised poses of keyframes `idx`."""
        self.T_WC[torch.as_tensor(idx, dtype=torch.long, device=self.device)] = poses.reshape(-1, 8)


class FactorGraph:
    """Array-level restatement of ``global_opt.FactorGraph`` (global_opt.py:14-268) on CUDA tensors.

    * ``add_factors`` (global_opt.py:49-138) takes the eight outputs of the symmetric matcher
      (or a callable producing them) instead of a model; the confidence combine, Q_conf gate and
      match fractions run in ``m3s_edge_confidence``; the edge filter is the reference's.
    * Edge rows are kept **two-way and pair-interleaved** (row 2e = i->j, row 2e+1 = j->i) in
      preallocated device tensors, so ``_prep_two_way_edges`` is a view, not five concatenations
      (global_opt.py:147-154), and both directions of a pair stay adjacent (the edge sharding of
      dist.py keeps them on one rank).  Gauss-Newton sums are order-independent up to fp64
      round-off, so this differs from the reference's [all i->j | all j->i] order by ~1e-16 relative.
    * ``solve_GN_rays`` / ``solve_GN_calib`` (global_opt.py:168-268) call the CUDA solvers and write
      the free poses back.  Keyframe ids are mapped to rows of the stacked arrays through the sorted
      unique list (identity when every keyframe is in the graph, the reference's implicit assumption).
    """

    def __init__(self, frames: KeyframeStore, K=None, img_size=None, cfg=None, capacity: int = 32,
                 calib_img_size_order: str = "reference"):
        self.frames = frames
        self.K = K
        self.img_size = img_size  # (H, W), needed by solve_GN_calib
        # The reference hands (H, W) = frame.img.shape[1:3] (global_opt.py:226) to gauss_newton_calib,
        # which unpacks it as (width, height) (gauss_newton_calib.py:75): on non-square images the
        # border gate compares u against H and v against W.  "reference" reproduces that (pinned by
        # tests/golden/mlx_graph.npz); "wh" passes the geometrically intended (W, H).
        if calib_img_size_order not in ("reference", "wh"):
            raise ValueError("calib_img_size_order must be 'reference' or 'wh'")
        self.calib_img_size_order = calib_img_size_order
        self.cfg = cfg or get_config()["local_opt"]
        self.device = frames.device
        self._e = 0  # undirected edges stored
        self._cap = 0
        self._grow(capacity)

    # -- storage ------------------------------------------------------------------------------
    def _grow(self, cap):
        n, dev = self.frames.n_points, self.device
        new = dict(ii2=torch.empty(2 * cap, dtype=torch.int32, device=dev),
                   jj2=torch.empty(2 * cap, dtype=torch.int32, device=dev),
                   idx2=torch.empty((2 * cap, n), dtype=torch.int32, device=dev),
                   valid2=torch.empty((2 * cap, n), dtype=torch.bool, device=dev),
                   Q2=torch.empty((2 * cap, n), dtype=torch.float32, device=dev))
        for k, t in new.items():
            if self._cap:
                t[:2 * self._e] = getattr(self, "_" + k)[:2 * self._e]
            setattr(self, "_" + k, t)
        self._cap = cap

    # the reference's attribute names, as views of the interleaved storage
    ii = property(lambda self: self._ii2[0:2 * self._e:2])
    jj = property(lambda self: self._jj2[0:2 * self._e:2])
    idx_ii2jj = property(lambda self: self._idx2[0:2 * self._e:2])
    idx_jj2ii = property(lambda self: self._idx2[1:2 * self._e:2])
    valid_match_j = property(lambda self: self._valid2[0:2 * self._e:2].unsqueeze(-1))
    valid_match_i = property(lambda self: self._valid2[1:2 * self._e:2].unsqueeze(-1))
    Q_ii2jj = property(lambda self: self._Q2[0:2 * self._e:2].unsqueeze(-1))
    Q_jj2ii = property(lambda self: self._Q2[1:2 * self._e:2].unsqueeze(-1))

    def __len__(self):
        return self._e

    # -- add_factors (global_opt.py:49-138) -----------------------------------------------------
    def add_factors(self, ii, jj, min_match_frac, match_out, is_reloc: bool = False) -> bool:
        if callable(match_out):
            match_out = match_out(ii, jj)
        idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qii, Qjj, Qji, Qij = match_out
        dev = self.device
        q_conf = self.cfg["Q_conf"]
        Qj, valid_j, frac_j = kernels.edge_confidence(_t(Qii, dev), _t(Qji, dev), _t(idx_i2j, dev),
                                                      _t(valid_match_j, dev), q_conf)
        Qi, valid_i, frac_i = kernels.edge_confidence(_t(Qjj, dev), _t(Qij, dev), _t(idx_j2i, dev),
                                                      _t(valid_match_i, dev), q_conf)
        ii_t = torch.as_tensor(list(ii), dtype=torch.int32, device=dev)
        jj_t = torch.as_tensor(list(jj), dtype=torch.int32, device=dev)
        invalid = torch.minimum(frac_j, frac_i) < min_match_frac
        invalid = (~(ii_t == (jj_t - 1))) & invalid          # consecutive keyframes always kept
        if bool(invalid.any()) and is_reloc:
            return False
        keep = torch.nonzero(~invalid).flatten()
        nk = int(keep.numel())
        if nk == 0:
            return False
        while self._e + nk > self._cap:
            self._grow(2 * self._cap)
        n = self.frames.n_points
        r = slice(2 * self._e, 2 * (self._e + nk))
        inter = lambda a, b: torch.stack([a, b], 1).reshape(2 * nk, *a.shape[1:])
        self._ii2[r] = inter(ii_t[keep], jj_t[keep])
        self._jj2[r] = inter(jj_t[keep], ii_t[keep])
        self._idx2[r] = inter(_t(idx_i2j, dev).reshape(-1, n)[keep].to(torch.int32),
                              _t(idx_j2i, dev).reshape(-1, n)[keep].to(torch.int32))
        # the stored masks are the matcher's (global_opt.py:131-134), not the Q-gated ones
        self._valid2[r] = inter(_t(valid_match_j, dev).reshape(-1, n)[keep].to(torch.bool),
                                _t(valid_match_i, dev).reshape(-1, n)[keep].to(torch.bool))
        self._Q2[r] = inter(Qj.reshape(-1, n)[keep], Qi.reshape(-1, n)[keep])
        self._e += nk
        return True

    # -- solver inputs ---------------------------------------------------------------------------
    def get_unique_kf_idx(self):
        """global_opt.py:140-145 (sorted unique keyframe ids of the graph)."""
        return torch.unique(torch.cat([self.ii, self.jj]))

    def _prep_two_way_edges(self):
        """global_opt.py:147-154, as views (pair-interleaved row order)."""
        m = 2 * self._e
        return self._ii2[:m], self._jj2[:m], self._idx2[:m], self._valid2[:m], self._Q2[:m]

    def _get_poses_points(self, unique_kf_idx):
        """global_opt.py:156-166.  No copy when the graph covers keyframes 0..K-1."""
        k = int(unique_kf_idx.numel())
        f = self.frames
        if k and int(unique_kf_idx[0]) == 0 and int(unique_kf_idx[-1]) == k - 1:
            return f.X[:k], f.T_WC[:k], f.Cavg[:k]
        sel = unique_kf_idx.long()
        return f.X[sel], f.T_WC[sel], f.Cavg[sel]

    def _rows(self, unique_kf_idx, ii, jj):
        k = int(unique_kf_idx.numel())
        if k and int(unique_kf_idx[0]) == 0 and int(unique_kf_idx[-1]) == k - 1:
            return ii, jj
        u = unique_kf_idx.to(torch.int32)
        return (torch.searchsorted(u, ii).to(torch.int32), torch.searchsorted(u, jj).to(torch.int32))

    # -- solves (global_opt.py:168-268) ------------------------------------------------------------
    def solve_GN_rays(self) -> None:
        pin = self.cfg["pin"]
        if self._e == 0:
            return
        uniq = self.get_unique_kf_idx()
        if uniq.numel() <= pin:
            return
        Xs, T, Cs = self._get_poses_points(uniq)
        ii, jj, idx, valid, Q = self._prep_two_way_edges()
        ii, jj = self._rows(uniq, ii, jj)
        out = solve_GN_rays(T, Xs, Cs, ii, jj, idx, valid, Q, self.cfg)
        self.frames.update_T_WCs(out[pin:], uniq[pin:])

    def solve_GN_calib(self) -> None:
        if self.K is None:
            raise ValueError("Intrinsic matrix K required for calibrated mode")
        if self.img_size is None:
            raise ValueError("img_size=(H, W) required for calibrated mode")
        pin = self.cfg["pin"]
        if self._e == 0:
            return
        uniq = self.get_unique_kf_idx()
        if uniq.numel() <= pin:
            return
        Xs, T, Cs = self._get_poses_points(uniq)
        Xs = kernels.constrain_points_to_ray(self.img_size, Xs, self.K)
        ii, jj, idx, valid, Q = self._prep_two_way_edges()
        ii, jj = self._rows(uniq, ii, jj)
        h, w = self.img_size
        size = (h, w) if self.calib_img_size_order == "reference" else (w, h)
        out = solve_GN_calib(T, Xs, Cs, self.K, ii, jj, idx, valid, Q, size, self.cfg)
        self.frames.update_T_WCs(out[pin:], uniq[pin:])


def _t(a, dev):
    return a.to(dev) if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a)).to(dev)
