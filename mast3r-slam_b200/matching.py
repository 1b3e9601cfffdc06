"""L2 matching host: drop-in for the hot-path entry points of the reference's ``matching.py``.

``match`` (matching.py:12-38), ``match_iterative_proj`` (:339-461) and ``prep_for_iter_proj``
(:134-175) keep the reference's names, argument order, tensor layouts and return values
(idx [B,H*W] int32, valid [B,H*W,1] bool).  The reference's tensors are MLX arrays; here they are
numpy arrays or torch CUDA tensors, and every arithmetic step runs in the sm_100a kernels
(one C-ABI call enqueues prep -> iter_proj -> refine -> tail; nothing is copied back in between,
unlike matching.py:374-377, 418-431).

``match_simple`` (matching.py:41-90) is the route ``match`` takes under the reference's *default*
configuration (config.py:65 ``use_simple=True``): identity / initial correspondences gated by the 3-D
distance.  It runs in ``m3s_match_simple`` (same arithmetic as the tail kernel).
"""
from __future__ import annotations

from . import kernels
from .config import get_config

prep_for_iter_proj = kernels.prep_for_iter_proj


def match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init=None, semantics=None):
    """matching.py:339-461.  X11, X21 [B,H,W,3]; D11, D21 [B,H,W,D]; idx_init [B,H*W] or None.
    ``semantics="metal"`` (or kernels.set_semantics): the steps the reference runs through its Metal
    shaders -- iter_proj (:393-405) and refine_matches (:411-416) -- take their Metal semantics; prep
    and tail are the same MLX code on either path."""
    cfg = get_config()["matching"]
    radius = cfg.get("refine_radius", 3)
    use_refine = cfg.get("use_refine", True) and radius > 0
    if kernels._is_metal(semantics):
        rays, pts, p0 = kernels.prep_for_iter_proj(X11, X21, idx_1_to_2_init)
        p, valid = kernels.iter_proj(rays, pts, p0, cfg["max_iter"], cfg["lambda_init"],
                                     cfg["convergence_thresh"], semantics="metal")
        b, h, w, _ = X11.shape
        if use_refine:
            d = D21.shape[-1]
            p = kernels.refine_matches(D11, D21.reshape(b, h * w, d), p, radius,
                                       cfg.get("refine_dilation", 2), semantics="metal")
        elif hasattr(p, "astype"):
            p = p.astype("int32")
        else:
            p = p.int()
        return kernels.match_tail(X11, X21, p, valid, cfg["dist_thresh"])
    return kernels.match_iterative_proj_device(
        X11, X21, D11, D21, idx_1_to_2_init, cfg["max_iter"], cfg["lambda_init"],
        cfg["convergence_thresh"], radius if use_refine else 0, cfg.get("refine_dilation", 2),
        cfg["dist_thresh"])


def match_simple(X11, X21, D11, D21, idx_1_to_2_init=None):
    """matching.py:41-90.  D11 / D21 are accepted and unused, as in the reference."""
    return kernels.match_simple(X11, X21, idx_1_to_2_init, get_config()["matching"]["dist_thresh"])


def match(X11, X21, D11, D21, idx_1_to_2_init=None):
    """matching.py:12-38: the config switch between match_simple (default) and match_iterative_proj."""
    if get_config().get("matching", {}).get("use_simple", True):
        return match_simple(X11, X21, D11, D21, idx_1_to_2_init)
    return match_iterative_proj(X11, X21, D11, D21, idx_1_to_2_init)


def match_symmetric(Xii, Xji, Xjj, Xij, Dii, Dji, Djj, Dij, idx_i2j_init=None, idx_j2i_init=None):
    """Both directions of a keyframe pair batch in ONE enqueue (what mast3r_match_symmetric,
    mast3r_utils.py:503-569, is meant to do once the decoder has produced the four pointmaps; the
    reference's version is an identity-match stub).  Inputs [B,H,W,3] / [B,H,W,D]; the i->j and
    j->i problems are stacked on the batch axis, so the pairs of a batch and the two directions all
    run in the same kernel launches.  Returns (idx_i2j, idx_j2i, valid_match_j, valid_match_i)."""
    import numpy as np
    import torch
    cat = (lambda a, b: torch.cat([a, b])) if isinstance(Xii, torch.Tensor) else \
          (lambda a, b: np.concatenate([np.asarray(a), np.asarray(b)]))
    init = None
    if idx_i2j_init is not None or idx_j2i_init is not None:
        if idx_i2j_init is None or idx_j2i_init is None:
            raise ValueError("match_symmetric: give both initialisations or neither")
        init = cat(idx_i2j_init, idx_j2i_init)
    idx, valid = match_iterative_proj(cat(Xii, Xjj), cat(Xji, Xij), cat(Dii, Djj), cat(Dji, Dij), init)
    b = Xii.shape[0]
    return idx[:b], idx[b:], valid[:b], valid[b:]
