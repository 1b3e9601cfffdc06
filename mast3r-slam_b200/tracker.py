"""L2 tracker host: the single-pose Sim3 solves of the reference's ``tracker.py``.

Mirrors /root/reference/src/mlx_mast3r_slam/tracker.py: ``FrameTracker._opt_pose_ray_dist_sim3``
(:255-313), ``_opt_pose_calib_sim3`` (:315-406), the confidence / mask block of ``track`` (:86-113)
and its keyframe-selection statistics (:151-164).  Poses are ``[8]`` rows ``(t, q_xyzw, s)`` like
``Sim3.data`` (liegroups/sim3.py:15).  numpy in -> numpy out, or torch CUDA tensors in -> torch out
(asynchronous on the current stream).  Everything numeric runs in ``csrc/track.cu`` behind the C
ABI (``m3s_track_gate``, ``m3s_track_opt_pose``); no fallback.

Out of scope (needs the MASt3R model / Frame objects): ``mast3r_match_fn``, ``update_pointmap``.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from .config import get_config
from .kernels import _dev, _device, _k4, _out, _ptr, _stream, _wants_numpy, _workspace
from ._lib import M3SError

__all__ = ["track_gate", "opt_pose_ray_dist_sim3", "opt_pose_calib_sim3", "FrameTracker"]


def _flat(x, dev, dtype, n=None, last=None):
    t = _dev(x, dtype, dev)
    t = t.reshape(-1, last) if last else t.reshape(-1)
    if n is not None and t.shape[0] != n:
        raise M3SError(f"tracker: expected {n} rows, got {t.shape[0]}")
    return t


def track_gate(Cf, Ck, Qff, Qkf, idx_f2k, valid_match_k, cfg=None):
    """tracker.py:86-113.  Returns (Qk [N,1] f32, valid_opt [N,1] bool, valid_kf [N,1] bool,
    (n_opt, n_kf)); the counts are read back (one 8-byte D2H) because ``track`` branches on them."""
    cfg = cfg or get_config()["tracking"]
    as_np = _wants_numpy(Cf, Ck, Qff, Qkf, idx_f2k, valid_match_k)
    dev = _device(Cf, Ck, Qff, Qkf, idx_f2k, valid_match_k)
    idx = _flat(idx_f2k, dev, torch.int32)
    n = idx.shape[0]
    cf, ck = _flat(Cf, dev, torch.float32, n), _flat(Ck, dev, torch.float32, n)
    qf, qk = _flat(Qff, dev, torch.float32, n), _flat(Qkf, dev, torch.float32, n)
    vm = _flat(valid_match_k, dev, torch.uint8, n)
    Qk = torch.empty(n, dtype=torch.float32, device=dev)
    v_opt = torch.empty(n, dtype=torch.uint8, device=dev)
    v_kf = torch.empty(n, dtype=torch.uint8, device=dev)
    counts = torch.empty(2, dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().m3s_track_gate(
            _ptr(cf), _ptr(ck), _ptr(qf), _ptr(qk), _ptr(idx), _ptr(vm), float(cfg["C_conf"]),
            float(cfg["Q_conf"]), _ptr(Qk), _ptr(v_opt), _ptr(v_kf), _ptr(counts), n, _stream(dev)),
            "track_gate")
    n_opt, n_kf = (int(v) for v in counts.cpu().tolist())
    return (_out(Qk[:, None], as_np), _out(v_opt[:, None].bool(), as_np),
            _out(v_kf[:, None].bool(), as_np), (n_opt, n_kf))


def _params(mode, cfg, K=None, img_size=None):
    p = _lib.TrackParams()
    p.mode = mode
    p.max_iters = int(cfg["max_iters"])
    p.sigma_ray, p.sigma_dist = float(cfg["sigma_ray"]), float(cfg["sigma_dist"])
    p.sigma_pixel, p.sigma_depth = float(cfg["sigma_pixel"]), float(cfg["sigma_depth"])
    p.huber = float(cfg["huber"])
    p.rel_error, p.delta_norm = float(cfg["rel_error"]), float(cfg["delta_norm"])
    p.pixel_border = int(cfg.get("pixel_border", 0))
    p.depth_eps = float(cfg.get("depth_eps", 0.0))
    if mode == _lib.TRACK_CALIB:
        p.img_h, p.img_w = int(img_size[0]), int(img_size[1])  # img_size = (h, w), tracker.py:100
        k4 = _k4(K)
        for i in range(4):
            p.K4[i] = k4[i]
    return p


def _opt_pose(mode, Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg, K=None, img_size=None, idx_f2k=None,
              return_stats=False):
    cfg = cfg or get_config()["tracking"]
    as_np = _wants_numpy(Xf, Xk, T_WCf, T_WCk, Qk, valid)
    dev = _device(Xf, Xk, T_WCf, T_WCk, Qk, valid)
    xk = _flat(Xk, dev, torch.float32, None, 3)
    n = xk.shape[0]
    xf = _flat(Xf, dev, torch.float32, n, 3)
    q = _flat(Qk, dev, torch.float32, n)
    v = _flat(valid, dev, torch.uint8, n)
    idx = None if idx_f2k is None else _flat(idx_f2k, dev, torch.int32, n)
    tf, tk = _flat(T_WCf, dev, torch.float32, 8), _flat(T_WCk, dev, torch.float32, 8)
    prm = _params(mode, cfg, K, img_size)
    out_f = torch.empty(8, dtype=torch.float32, device=dev)
    out_r = torch.empty(8, dtype=torch.float32, device=dev)
    stats = torch.empty(4, dtype=torch.float64, device=dev)
    lib = _lib.load()
    with torch.cuda.device(dev):
        nbytes = lib.m3s_track_workspace_bytes(n)
        ws = _workspace(dev, nbytes)
        _lib.check(lib.m3s_track_opt_pose(
            C.byref(prm), _ptr(xf), _ptr(idx) if idx is not None else None, _ptr(xk), _ptr(q), _ptr(v),
            n, _ptr(tf), _ptr(tk), _ptr(out_f), _ptr(out_r), _ptr(stats), _ptr(ws), ws.numel(),
            _stream(dev)), "track_opt_pose")
    res = (_out(out_f, as_np), _out(out_r, as_np))
    if return_stats:
        s = stats.cpu().tolist()
        if s[3] != 0.0:
            raise M3SError("track_opt_pose: normal equations not positive definite / non-finite")
        return res + ({"iterations": int(s[0]), "cost": s[1], "delta_norm": s[2]},)
    return res


def opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, idx_f2k=None, return_stats=False):
    """tracker.py:255-313.  Xf [N,3] (already gathered ``Xf[idx_f2k]`` as in ``_get_points_poses``
    :214, or raw with ``idx_f2k`` given), Xk [N,3], Qk [N,1], valid [N,1] -> (T_WCf [8], T_CkCf [8])."""
    return _opt_pose(_lib.TRACK_RAY_DIST, Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg, idx_f2k=idx_f2k,
                     return_stats=return_stats)


def opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, K, img_size, cfg=None, idx_f2k=None,
                        return_stats=False):
    """tracker.py:315-406.  ``meas_k`` / ``valid_meas_k`` (:199-212) are derived inside the kernel
    from Xk and the pixel grid; Xf / Xk are expected ray-constrained (``constrain_points_to_ray``,
    :196-197).  img_size = (h, w)."""
    return _opt_pose(_lib.TRACK_CALIB, Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg, K, img_size, idx_f2k,
                     return_stats)


class FrameTracker:
    """Array-level restatement of ``FrameTracker`` (tracker.py:26-180) without Frame / model
    objects: ``track_arrays`` takes what ``mast3r_match_fn`` returned plus the two frames' canonical
    points, confidences and poses, and returns the new frame pose and the keyframe decision."""

    def __init__(self, cfg=None):
        self.cfg = dict(cfg or get_config()["tracking"])
        self.idx_f2k = None

    def reset_idx_f2k(self) -> None:
        self.idx_f2k = None

    def _opt_pose_ray_dist_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid):
        return opt_pose_ray_dist_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, self.cfg)

    def _opt_pose_calib_sim3(self, Xf, Xk, T_WCf, T_WCk, Qk, valid, K, img_size):
        return opt_pose_calib_sim3(Xf, Xk, T_WCf, T_WCk, Qk, valid, K, img_size, self.cfg)

    def track_arrays(self, idx_f2k, valid_match_k, Xf_canon, Cf, Qff, Xk_canon, Ck, Qkf, T_WCf, T_WCk,
                     K=None, img_size=None):
        """tracker.py:51-180 on arrays.  Returns dict(new_kf, try_reloc, T_WCf, T_CkCf, match_frac)."""
        cfg = self.cfg
        self.idx_f2k = idx_f2k
        Qk, v_opt, v_kf, (n_opt, n_kf) = track_gate(Cf, Ck, Qff, Qkf, idx_f2k, valid_match_k, cfg)
        n = int(np.prod(v_opt.shape))
        if n == 0 or n_opt / n < cfg["min_match_frac"]:  # tracker.py:116-119
            return {"new_kf": False, "try_reloc": True, "T_WCf": T_WCf, "T_CkCf": None,
                    "match_frac": (n_opt / n if n else 0.0)}
        use_calib = bool(cfg.get("use_calib", False)) and K is not None
        if use_calib:
            from .kernels import constrain_points_to_ray
            shp = (1, n, 3)
            Xf_c = constrain_points_to_ray(img_size, Xf_canon.reshape(shp), K).reshape(n, 3)
            Xk_c = constrain_points_to_ray(img_size, Xk_canon.reshape(shp), K).reshape(n, 3)
            T_new, T_rel = opt_pose_calib_sim3(Xf_c, Xk_c, T_WCf, T_WCk, Qk, v_opt, K, img_size, cfg,
                                               idx_f2k=idx_f2k)
        else:
            T_new, T_rel = opt_pose_ray_dist_sim3(Xf_canon, Xk_canon, T_WCf, T_WCk, Qk, v_opt, cfg,
                                                  idx_f2k=idx_f2k)
        # keyframe selection (tracker.py:151-164)
        idx_t = torch.as_tensor(np.asarray(idx_f2k) if not isinstance(idx_f2k, torch.Tensor) else idx_f2k)
        vm_t = torch.as_tensor(np.asarray(valid_match_k) if not isinstance(valid_match_k, torch.Tensor)
                               else valid_match_k).reshape(-1).bool()
        unique_frac_f = int(torch.unique(idx_t.reshape(-1)[vm_t.to(idx_t.device)]).numel()) / n
        new_kf = min(n_kf / n, unique_frac_f) < cfg["match_frac_thresh"]
        if new_kf:
            self.reset_idx_f2k()
        return {"new_kf": bool(new_kf), "try_reloc": False, "T_WCf": T_new, "T_CkCf": T_rel,
                "match_frac": n_opt / n}
