"""Hot-path configuration defaults (the part of the reference's config that reaches the kernels).

Values are those of /root/reference/src/mlx_mast3r_slam/config.py:55-114 (DEFAULT_CONFIG), plus the
three keys match_iterative_proj reads with inline defaults (matching.py:405-407: refine_radius=3,
refine_dilation=2, use_refine=True).  No YAML/inherit machinery: out of scope.
"""
from __future__ import annotations

import copy

DEFAULT_CONFIG = {
    "matching": {
        "use_simple": True,          # config.py:65 -- the reference's default bypasses the hot path
        "use_metal": True,
        "max_iter": 10,
        "lambda_init": 1e-8,
        "convergence_thresh": 1e-6,
        "dist_thresh": 0.1,
        "radius": 3,
        "dilation_max": 0,
        "refine_radius": 3,          # matching.py:405
        "refine_dilation": 2,        # matching.py:406
        "use_refine": True,          # matching.py:407
    },
    "tracking": {                    # config.py:74-91
        "min_match_frac": 0.05, "C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3, "delta_norm": 1e-3,
        "max_iters": 10, "huber": 1.345, "sigma_ray": 0.003, "sigma_dist": 10.0, "sigma_pixel": 1.0,
        "sigma_depth": 10.0, "pixel_border": 0, "depth_eps": 0.0, "match_frac_thresh": 0.333,
        "use_calib": False,
    },
    "local_opt": {
        "pin": 1, "max_iters": 10, "C_conf": 0.0, "Q_conf": 1.5, "sigma_ray": 0.003,
        "sigma_dist": 10.0, "sigma_pixel": 1.0, "sigma_depth": 10.0, "pixel_border": 0,
        "depth_eps": 0.0, "delta_norm": 1e-3,
    },
}

_config = copy.deepcopy(DEFAULT_CONFIG)


def get_config() -> dict:
    """config.py:117-121: the active configuration (defaults unless set_config was called)."""
    return _config


def set_config(cfg: dict) -> None:
    """Shallow-merge `cfg` sections over the defaults."""
    global _config
    merged = copy.deepcopy(DEFAULT_CONFIG)
    for k, v in cfg.items():
        if isinstance(v, dict) and k in merged:
            merged[k].update(v)
        else:
            merged[k] = v
    _config = merged
