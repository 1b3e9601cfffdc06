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
