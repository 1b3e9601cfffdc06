"""L1 operator API: drop-in for the reference's ``backends/mpsgraph/kernels.py``.

Same function names, argument order, defaults and return conventions as
/root/reference/src/mlx_mast3r_slam/backends/mpsgraph/kernels.py (iter_proj :107,
gauss_newton_rays :262, gauss_newton_calib :325, gauss_newton_points :396, refine_matches :463,
is_available :27).  Semantics are those of the reference's numpy ("CPU") path, which is what its
own dispatcher lands on when the Metal runner is unavailable.

Inputs may be numpy arrays (reference convention: numpy in -> numpy out, host<->device copies
happen inside the call) or torch CUDA tensors (zero-copy, torch tensors out, asynchronous on the
current stream).  All arithmetic runs in the hand-written sm_100a kernels behind the C ABI
(include/m3s.h); torch is used only for device memory and streams.  ``use_metal`` is accepted for
signature compatibility and ignored.  Unlike the reference there is NO silent fallback: errors
raise ``M3SError``.
"""
from __future__ import annotations

import ctypes as C
import os
from collections import OrderedDict

import numpy as np
import torch

from . import _lib
from ._lib import GNParams, M3SError  # noqa: F401

__all__ = ["is_available", "iter_proj", "refine_matches", "gauss_newton_rays",
           "gauss_newton_points", "gauss_newton_calib", "match_iterative_proj_device", "match_iterative_proj_host",
           "prep_for_iter_proj", "match_tail", "GNPlan", "constrain_points_to_ray", "edge_confidence"]

_workspaces: dict = {}


def is_available() -> bool:
    """kernels.py:27-29.  True when the CUDA extension is built and a device is visible."""
    try:
        return _lib.load().m3s_device_count() > 0 and torch.cuda.is_available()
    except Exception:
        return False


# ---------------------------------------------------------------------------------------------
# tensor plumbing
# ---------------------------------------------------------------------------------------------
def _device(*xs) -> torch.device:
    for x in xs:
        if isinstance(x, torch.Tensor) and x.is_cuda:
            return x.device
    _lib.require_device()
    return torch.device("cuda", torch.cuda.current_device())


def _dev(x, dtype, device):
    """numpy / torch -> contiguous CUDA tensor of `dtype` (no copy when already there)."""
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=dtype, non_blocking=True).contiguous()
    a = np.asarray(x)
    np_dtype = {torch.float32: np.float32, torch.int32: np.int32, torch.uint8: np.uint8,
                torch.bool: np.bool_}[dtype]
    if a.dtype != np_dtype:
        a = a.astype(np_dtype)
    a = np.ascontiguousarray(a)
    t = torch.from_numpy(a)
    return t.to(device=device, non_blocking=True)


def _ptr(t) -> C.c_void_p:
    return C.c_void_p(t.data_ptr())


def _stream(device) -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _workspace(device, nbytes: int) -> torch.Tensor:
    key = (device.index, torch.cuda.current_stream(device).cuda_stream)
    ws = _workspaces.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(int(nbytes), 1024), dtype=torch.uint8, device=device)
        _workspaces[key] = ws
    return ws


def _wants_numpy(*xs) -> bool:
    return not any(isinstance(x, torch.Tensor) for x in xs)


def _out(t, as_numpy, np_dtype=None):
    if not as_numpy:
        return t
    a = t.cpu().numpy()
    return a.astype(np_dtype) if np_dtype is not None and a.dtype != np_dtype else a


# ---------------------------------------------------------------------------------------------
# factor-graph host steps (global_opt.py:86-104, geometry.py:273-302)
# ---------------------------------------------------------------------------------------------
def _k4(K):
    k = np.asarray(K.detach().cpu().numpy() if isinstance(K, torch.Tensor) else K, dtype=np.float32)
    if k.shape == (3, 3):
        k = np.array([k[0, 0], k[1, 1], k[0, 2], k[1, 2]], dtype=np.float32)
    if k.shape != (4,):
        raise M3SError("K must be a 3x3 matrix or (fx, fy, cx, cy)")
    return (C.c_float * 4)(*[float(v) for v in k])


def constrain_points_to_ray(img_size, Xs, K):
    """geometry.py:273-302: re-project every point onto its own pixel ray at its own depth.
    img_size = (H, W); Xs [B,H*W,3]; K 3x3 or (fx,fy,cx,cy)."""
    as_np = _wants_numpy(Xs)
    dev = _device(Xs)
    h, w = int(img_size[0]), int(img_size[1])
    x = _dev(Xs, torch.float32, dev)
    if x.dim() != 3 or x.shape[1] != h * w or x.shape[2] != 3:
        raise M3SError("constrain_points_to_ray: Xs must be [B,H*W,3]")
    out = torch.empty_like(x)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().m3s_constrain_points_to_ray(_ptr(x), _k4(K), _ptr(out), x.shape[0], h, w,
                                                         _stream(dev)), "constrain_points_to_ray")
    return _out(out, as_np)


def edge_confidence(Qii, Qji, idx, valid_match, Q_conf):
    """global_opt.py:86-104 -> (Qj [E,N,1] f32, valid [E,N,1] bool, match_frac [E] f32)."""
    as_np = _wants_numpy(Qii, Qji, idx, valid_match)
    dev = _device(Qii, Qji, idx, valid_match)
    qi = _dev(Qii, torch.float32, dev)
    e, n = qi.shape[0], qi.shape[1]
    qi = qi.reshape(e, n)
    qj = _dev(Qji, torch.float32, dev).reshape(e, n)
    ix = _dev(idx, torch.int32, dev).reshape(e, n)
    vm = _dev(valid_match, torch.bool, dev).reshape(e, n).view(torch.uint8)
    q_out = torch.empty((e, n), dtype=torch.float32, device=dev)
    v_out = torch.empty((e, n), dtype=torch.uint8, device=dev)
    cnt = torch.empty((e,), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        _lib.check(_lib.load().m3s_edge_confidence(_ptr(qi), _ptr(qj), _ptr(ix), _ptr(vm), float(Q_conf),
                                                 _ptr(q_out), _ptr(v_out), _ptr(cnt), e, n, _stream(dev)),
                   "edge_confidence")
    frac = cnt.to(torch.float32) / float(max(n, 1))
    return (_out(q_out.reshape(e, n, 1), as_np), _out(v_out.view(torch.bool).reshape(e, n, 1), as_np),
            _out(frac, as_np))


# ---------------------------------------------------------------------------------------------
# matching
# ---------------------------------------------------------------------------------------------
def prep_for_iter_proj(X11, X21, idx_1_to_2_init=None):
    """matching.py:134-175 -> (rays_with_grad [B,H,W,9], pts3d_norm [B,N,3], p_init [B,N,2])."""
    as_np = _wants_numpy(X11, X21)
    dev = _device(X11, X21)
    with torch.cuda.device(dev):
        X11d, X21d = _dev(X11, torch.float32, dev), _dev(X21, torch.float32, dev)
        b, h, w, _ = X11d.shape
        idx = None if idx_1_to_2_init is None else _dev(idx_1_to_2_init, torch.int32, dev)
        rays = torch.empty((b, h, w, 9), dtype=torch.float32, device=dev)
        pts = torch.empty((b, h * w, 3), dtype=torch.float32, device=dev)
        p0 = torch.empty((b, h * w, 2), dtype=torch.float32, device=dev)
        _lib.check(_lib.load().m3s_prep_iter_proj(
            _ptr(X11d), _ptr(X21d), _ptr(idx) if idx is not None else None, _ptr(rays), _ptr(pts),
            _ptr(p0), b, h, w, _stream(dev)), "prep_iter_proj")
    return _out(rays, as_np), _out(pts, as_np), _out(p0, as_np)


_semantics = os.environ.get("M3S_SEMANTICS", "numpy")


def set_semantics(semantics: str):
    """Which of the reference's two code paths every later call reproduces:
    "numpy" (default; its numpy path, the parity bar) or "metal" (its Metal shaders: per-point early
    exit and fp32 blend in iter_proj, in-place coarse-to-fine cascade in refine_matches, fp32
    Gauss-Newton with dt = tau; include/m3s.h M3S_SEM_*).  Every operator also takes
    ``semantics=...`` for one call.  ``use_metal`` does NOT select it: the reference's callers pass
    use_metal=True everywhere and fall back to numpy without a GPU, which is the behaviour the
    drop-in mirrors."""
    global _semantics
    if semantics not in ("numpy", "metal"):
        raise ValueError(f"unknown semantics {semantics!r}")
    _semantics = semantics


def _is_metal(semantics) -> bool:
    s = _semantics if semantics is None else semantics
    if s not in ("numpy", "metal"):
        raise ValueError(f"unknown semantics {s!r}")
    return s == "metal"


def iter_proj(rays_with_grad, pts3d_norm, p_init, max_iter: int = 10, lambda_init: float = 1e-8,
              convergence_thresh: float = 1e-6, use_metal: bool = True, semantics=None):
    """kernels.py:107-148 -> (p_final [B,N,2] f32, valid [B,N] bool)."""
    as_np = _wants_numpy(rays_with_grad, pts3d_norm, p_init)
    dev = _device(rays_with_grad, pts3d_norm, p_init)
    with torch.cuda.device(dev):
        rays = _dev(rays_with_grad, torch.float32, dev)
        pts = _dev(pts3d_norm, torch.float32, dev)
        p0 = _dev(p_init, torch.float32, dev)
        b, h, w, c = rays.shape
        if c != 9:
            raise M3SError("iter_proj: rays_with_grad must be [B,H,W,9]")
        n = pts.shape[1]
        p = torch.empty((b, n, 2), dtype=torch.float32, device=dev)
        valid = torch.empty((b, n), dtype=torch.uint8, device=dev)
        if b * n == 0:
            return _out(p, as_np), _out(valid.view(torch.bool), as_np)
        lib = _lib.load()
        if _is_metal(semantics):
            _lib.check(lib.m3s_iter_proj_metal(_ptr(rays), _ptr(pts), _ptr(p0), _ptr(p), _ptr(valid), b, h,
                                               w, n, int(max_iter), float(lambda_init),
                                               float(convergence_thresh), _stream(dev)), "iter_proj_metal")
            return _out(p, as_np), _out(valid.view(torch.bool), as_np)
        nb = lib.m3s_iter_proj_workspace_bytes(int(max_iter))
        ws = _workspace(dev, nb)
        _lib.check(lib.m3s_iter_proj(_ptr(rays), _ptr(pts), _ptr(p0), _ptr(p), _ptr(valid), b, h, w,
                                     n, int(max_iter), float(lambda_init),
                                     float(convergence_thresh), _ptr(ws), ws.numel(), _stream(dev)),
                   "iter_proj")
        valid = valid.view(torch.bool)
    return _out(p, as_np), _out(valid, as_np)


def refine_matches(D11, D21, p1, radius: int = 3, dilation_max: int = 0, use_metal: bool = True,
                   semantics=None):
    """kernels.py:463-493 -> p1_refined [B,N,2] int32.  p1 may be integer or float (truncated)."""
    as_np = _wants_numpy(D11, D21, p1)
    dev = _device(D11, D21, p1)
    with torch.cuda.device(dev):
        D11d, D21d = _dev(D11, torch.float32, dev), _dev(D21, torch.float32, dev)
        b, h, w, d = D11d.shape
        if not isinstance(p1, torch.Tensor):
            p1 = np.asarray(p1)
            if p1.dtype != np.float32 and p1.dtype != np.int32:
                p1 = p1.astype(np.int32)  # int(p1[...]) truncation of kernels.py:508,515
        is_float = (p1.dtype.is_floating_point if isinstance(p1, torch.Tensor)
                    else np.issubdtype(p1.dtype, np.floating))
        p1d = _dev(p1, torch.float32 if is_float else torch.int32, dev)
        n = p1d.shape[1]
        out = torch.empty((b, n, 2), dtype=torch.int32, device=dev)
        if b * n == 0:
            return _out(out, as_np)
        if _is_metal(semantics):
            if is_float:
                p1d = p1d.to(torch.int32)   # p1.astype(np.int32), refine_metal_runner.py:120
            fn = _lib.load().m3s_refine_matches_metal
        else:
            fn = _lib.load().m3s_refine_matches_f32 if is_float else _lib.load().m3s_refine_matches
        _lib.check(fn(_ptr(D11d), _ptr(D21d), _ptr(p1d), _ptr(out), b, h, w, d, n, int(radius),
                      int(dilation_max), _stream(dev)), "refine_matches")
    return _out(out, as_np)


def match_tail(X11, X21, p1_int, valid_proj, dist_thresh: float = 0.1):
    """matching.py:438-461 -> (idx [B,N] i32, valid [B,N,1] bool)."""
    as_np = _wants_numpy(X11, X21, p1_int, valid_proj)
    dev = _device(X11, X21, p1_int, valid_proj)
    with torch.cuda.device(dev):
        X11d, X21d = _dev(X11, torch.float32, dev), _dev(X21, torch.float32, dev)
        b, h, w, _ = X11d.shape
        p1d = _dev(p1_int, torch.int32, dev)
        vp = valid_proj
        vp = (vp.to(dev).to(torch.uint8) if isinstance(vp, torch.Tensor)
              else _dev(np.asarray(vp).astype(np.uint8), torch.uint8, dev)).contiguous()
        idx = torch.empty((b, h * w), dtype=torch.int32, device=dev)
        valid = torch.empty((b, h * w), dtype=torch.uint8, device=dev)
        _lib.check(_lib.load().m3s_match_tail(_ptr(X11d), _ptr(X21d), _ptr(p1d), _ptr(vp), _ptr(idx),
                                              _ptr(valid), b, h, w, float(dist_thresh),
                                              _stream(dev)), "match_tail")
        valid = valid.view(torch.bool)[:, :, None]
    return _out(idx, as_np), _out(valid, as_np)


def match_simple(X11, X21, idx_init=None, dist_thresh: float = 0.1):
    """matching.py:41-90 (the reference's default route) -> (idx [B,N] i32, valid [B,N,1] bool)."""
    as_np = _wants_numpy(X11, X21)
    dev = _device(X11, X21, idx_init)
    with torch.cuda.device(dev):
        X11d, X21d = _dev(X11, torch.float32, dev), _dev(X21, torch.float32, dev)
        b, h, w = X21d.shape[:3]
        idx0 = None if idx_init is None else _dev(idx_init, torch.int32, dev).reshape(b, h * w)
        idx = torch.empty((b, h * w), dtype=torch.int32, device=dev)
        valid = torch.empty((b, h * w), dtype=torch.uint8, device=dev)
        _lib.check(_lib.load().m3s_match_simple(
            _ptr(X11d), _ptr(X21d), _ptr(idx0) if idx0 is not None else None, _ptr(idx), _ptr(valid),
            b, h, w, float(dist_thresh), _stream(dev)), "match_simple")
        valid = valid.view(torch.bool)[:, :, None]
    return _out(idx, as_np), _out(valid, as_np)


def match_iterative_proj_device(X11, X21, D11, D21, idx_init=None, max_iter=10, lambda_init=1e-8,
                                convergence_thresh=1e-6, radius=3, dilation_max=2, dist_thresh=0.1):
    """Whole match_iterative_proj (matching.py:339-461) as ONE C-ABI call on device tensors.
    Returns (idx [B,N] i32, valid [B,N,1] bool) as CUDA tensors (or numpy when given numpy)."""
    if not any(isinstance(x, torch.Tensor) and x.is_cuda for x in (X11, X21, D11, D21)):
        return match_iterative_proj_host(X11, X21, D11, D21, idx_init, max_iter, lambda_init,
                                         convergence_thresh, radius, dilation_max, dist_thresh)
    as_np = _wants_numpy(X11, X21, D11, D21)
    dev = _device(X11, X21, D11, D21)
    with torch.cuda.device(dev):
        X11d, X21d = _dev(X11, torch.float32, dev), _dev(X21, torch.float32, dev)
        b, h, w, _ = X11d.shape
        D11d = _dev(D11, torch.float32, dev).reshape(b, h, w, -1)
        D21d = _dev(D21, torch.float32, dev).reshape(b, h * w, -1)
        d = D11d.shape[-1]
        idx0 = None if idx_init is None else _dev(idx_init, torch.int32, dev)
        idx = torch.empty((b, h * w), dtype=torch.int32, device=dev)
        valid = torch.empty((b, h * w), dtype=torch.uint8, device=dev)
        lib = _lib.load()
        nb = lib.m3s_match_workspace_bytes(b, h, w, int(max_iter))
        ws = _workspace(dev, nb)
        _lib.check(lib.m3s_match_iterative_proj(
            _ptr(X11d), _ptr(X21d), _ptr(D11d), _ptr(D21d), _ptr(idx0) if idx0 is not None else None,
            _ptr(idx), _ptr(valid), b, h, w, d, int(max_iter), float(lambda_init),
            float(convergence_thresh), int(radius), int(dilation_max), float(dist_thresh), _ptr(ws),
            ws.numel(), _stream(dev)), "match_iterative_proj")
        valid = valid.view(torch.bool)[:, :, None]
    return _out(idx, as_np), _out(valid, as_np)


def _host_f32(x):
    """host array -> (C-contiguous float32 numpy view, keep-alive object); no copy when possible"""
    if isinstance(x, torch.Tensor):
        t = x.detach().to(torch.float32).contiguous()
        return t.numpy(), t
    a = np.ascontiguousarray(np.asarray(x), dtype=np.float32)
    return a, a


def match_iterative_proj_host(X11, X21, D11, D21, idx_init=None, max_iter=10, lambda_init=1e-8,
                              convergence_thresh=1e-6, radius=3, dilation_max=2, dist_thresh=0.1,
                              chunk_pairs=4, out=None, device=None):
    """match_iterative_proj (matching.py:339-461) on HOST arrays (numpy or CPU torch tensors, the
    reference's calling convention) through the chunk-pipelined C-ABI entry point
    m3s_match_iterative_proj_host: upload, kernels and download of consecutive chunks of
    ``chunk_pairs`` pairs overlap on three streams.  Page-locked inputs (torch ``pin_memory``)
    make the copies asynchronous; ``out = (idx int32 [B,N], valid uint8/bool [B,N])`` lets the
    caller provide (page-locked) result arrays.  Returns (idx [B,N] i32, valid [B,N,1] bool) as
    numpy arrays (CPU torch tensors when the inputs were torch tensors)."""
    _lib.require_device()
    as_torch = any(isinstance(x, torch.Tensor) for x in (X11, X21, D11, D21))
    x11, k1 = _host_f32(X11)
    x21, k2 = _host_f32(X21)
    b, h, w, _ = x11.shape
    use_desc = radius > 0
    d11 = d21 = None
    d = 0
    if use_desc or D11 is not None:
        d11, k3 = _host_f32(D11)
        d21, k4 = _host_f32(D21)
        d = d11.shape[-1]
    init = None
    if idx_init is not None:
        init = (idx_init.detach().to(torch.int32).contiguous().numpy() if isinstance(idx_init, torch.Tensor)
                else np.ascontiguousarray(np.asarray(idx_init), dtype=np.int32))
    if out is not None:
        idx_o, valid_o = out
        idx_a = idx_o.numpy() if isinstance(idx_o, torch.Tensor) else idx_o
        valid_a = valid_o.numpy() if isinstance(valid_o, torch.Tensor) else valid_o
        if idx_a.dtype != np.int32 or valid_a.dtype not in (np.uint8, np.bool_) or \
                idx_a.size != b * h * w or valid_a.size != b * h * w or \
                not idx_a.flags.c_contiguous or not valid_a.flags.c_contiguous:
            raise M3SError("match_iterative_proj_host: out must be C-contiguous (int32 [B,N], uint8/bool [B,N])")
    else:
        idx_a = np.empty((b, h * w), np.int32)
        valid_a = np.empty((b, h * w), np.bool_)
    dev_index = (torch.cuda.current_device() if device is None
                 else torch.device(device).index or 0)
    vp = lambda a: None if a is None else C.c_void_p(a.ctypes.data)  # noqa: E731
    _lib.check(_lib.load().m3s_match_iterative_proj_host(
        vp(x11), vp(x21), vp(d11), vp(d21), vp(init), vp(idx_a), vp(valid_a), b, h, w, d,
        int(max_iter), float(lambda_init), float(convergence_thresh), int(radius), int(dilation_max),
        float(dist_thresh), int(chunk_pairs), int(dev_index)), "match_iterative_proj_host")
    idx_r = idx_a.reshape(b, h * w)
    valid_r = valid_a.reshape(b, h * w).view(np.bool_)[:, :, None]
    if as_torch:
        return torch.from_numpy(idx_r), torch.from_numpy(valid_r)
    return idx_r, valid_r


# ---------------------------------------------------------------------------------------------
# Gauss-Newton
# ---------------------------------------------------------------------------------------------
class GNPlan:
    """Owner of a ``m3s_gn_plan`` (graph topology + device workspaces); see include/m3s.h."""

    def __init__(self, ii, jj, num_kf: int, num_pts: int, pin: int, device: torch.device):
        self.ii = np.ascontiguousarray(np.asarray(ii, dtype=np.int32).reshape(-1))
        self.jj = np.ascontiguousarray(np.asarray(jj, dtype=np.int32).reshape(-1))
        if self.ii.shape != self.jj.shape:
            raise M3SError("ii and jj must have the same length")
        self.E, self.num_kf, self.N, self.pin, self.device = len(self.ii), num_kf, num_pts, pin, device
        lib = _lib.load()
        _lib.require_device()
        self._lib = lib
        self.handle = lib.m3s_gn_plan_create(
            self.ii.ctypes.data_as(C.c_void_p), self.jj.ctypes.data_as(C.c_void_p), self.E,
            int(num_kf), int(num_pts), int(pin), int(device.index or 0))
        if not self.handle:
            raise M3SError("gn_plan_create failed: " + lib.m3s_last_error().decode())
        self.F = lib.m3s_gn_plan_num_free(self.handle)
        self.dim = lib.m3s_gn_plan_dim(self.handle)
        self.ld = lib.m3s_gn_plan_ld(self.handle)

    def __del__(self):
        h, self.handle = getattr(self, "handle", None), None
        if h:
            try:
                self._lib.m3s_gn_plan_destroy(h)
            except Exception:
                pass

    def _view(self, ptr, count):
        """Zero-copy torch view (float64) of a plan-owned device buffer."""
        if not ptr or count == 0:
            return torch.zeros(0, dtype=torch.float64, device=self.device)
        iface = {"shape": (int(count),), "typestr": "<f8", "data": (int(ptr), False), "version": 3,
                 "strides": None}
        holder = type("_Buf", (), {"__cuda_array_interface__": iface})()
        t = torch.as_tensor(holder, device=self.device)
        t._m3s_plan = self  # keep the plan alive while the view lives
        return t

    def edge_blocks(self) -> torch.Tensor:
        n = C.c_int64(0)
        p = self._lib.m3s_gn_plan_edge_blocks(self.handle, C.byref(n))
        return self._view(p, n.value).view(self.E, _lib.GN_EDGE_STRIDE) if self.E else self._view(0, 0)

    def H(self) -> torch.Tensor:
        """[dim, dim] strided view of the assembled H (or its Cholesky factor after a solve)."""
        if self.dim == 0:
            return torch.zeros((0, 0), dtype=torch.float64, device=self.device)
        full = self._view(self._lib.m3s_gn_plan_H(self.handle), (self.ld + 1) * self.ld)
        return full.view(self.ld + 1, self.ld)[: self.dim, : self.dim]

    def g(self) -> torch.Tensor:
        return self._view(self._lib.m3s_gn_plan_g(self.handle), self.dim)

    def poses(self) -> torch.Tensor:
        """[num_kf, 8] fp64 pose state of the solver (before the final float32 cast)."""
        return self._view(self._lib.m3s_gn_plan_poses(self.handle), 8 * self.num_kf).view(-1, 8)

    def factor_path(self) -> int:
        """0 = one launch per block step, 1 = persistent chain + workers, 2 = dataflow over
        shared-memory-resident tiles, 3 = sparse front + path 2 on the dense tail
        (m3s_gn_plan_factor_path; decided from M3S_GN_CHOL / M3S_GN_SPARSE, the graph and what is
        co-resident on the device)"""
        return int(_lib.load().m3s_gn_plan_factor_path(self.handle))

    def sparse_info(self) -> dict:
        """levels / eliminated keyframes / tail keyframes / padded tail dimension of the sparse front"""
        out = (C.c_int32 * 4)()
        _lib.check(_lib.load().m3s_gn_plan_sparse_info(self.handle, out))
        return {"levels": int(out[0]), "eliminated": int(out[1]), "tail": int(out[2]), "tail_np": int(out[3])}

    def dx(self) -> torch.Tensor:
        return self._view(self._lib.m3s_gn_plan_dx(self.handle), self.dim)

    def status(self):
        st = (C.c_int32 * 4)()
        nrm = C.c_double(0)
        _lib.check(self._lib.m3s_gn_status(self.handle, st, C.byref(nrm), _stream(self.device)))
        return dict(applied=st[0], stopped=bool(st[1]), failed=bool(st[2]), attempted=st[3],
                    last_dx_norm=nrm.value, exchange_timeout=(st[2] == 2))


_plan_cache: "OrderedDict[tuple, GNPlan]" = OrderedDict()
_PLAN_CACHE_MAX = 8


def get_plan(ii, jj, num_kf, num_pts, pin, device) -> GNPlan:
    ii = np.ascontiguousarray(np.asarray(ii, dtype=np.int32).reshape(-1))
    jj = np.ascontiguousarray(np.asarray(jj, dtype=np.int32).reshape(-1))
    key = (ii.tobytes(), jj.tobytes(), int(num_kf), int(num_pts), int(pin), device.index)
    plan = _plan_cache.get(key)
    if plan is None:
        plan = GNPlan(ii, jj, num_kf, num_pts, pin, device)
        _plan_cache[key] = plan
        while len(_plan_cache) > _PLAN_CACHE_MAX:
            _plan_cache.popitem(last=False)
    else:
        _plan_cache.move_to_end(key)
    return plan


_gn_precision = None  # None = library default (mixed unless M3S_GN_PRECISION=f64)


def set_gn_precision(precision):
    """Per-point arithmetic of the Gauss-Newton accumulate pass for every later call that does not
    name one itself: None (library default), "f64" (the reference's numpy arithmetic) or "mixed"
    (fp32 per-point products, fp64 reductions / H / solve / poses); see include/m3s.h."""
    global _gn_precision
    if precision not in (None, "default", "f64", "mixed"):
        raise ValueError(f"unknown GN precision {precision!r}")
    _gn_precision = None if precision == "default" else precision


def make_params(mode, max_iter, pin, delta_thresh, C_thresh, Q_thresh, sigma_a, sigma_b=1.0,
                K=None, img_size=(0, 0), pixel_border=0, z_eps=0.0, precision=None,
                semantics=None) -> GNParams:
    """``precision``: None / "default", "f64" or "mixed" (include/m3s.h, M3S_GN_PREC_*);
    ``semantics``: None (module setting), "numpy" or "metal" (M3S_SEM_*)."""
    p = GNParams()
    p.semantics = 1 if _is_metal(semantics) else 0
    if precision is None:
        precision = _gn_precision
    p.precision = {None: _lib.GN_PREC_DEFAULT, "default": _lib.GN_PREC_DEFAULT,
                   "f64": _lib.GN_PREC_F64, "mixed": _lib.GN_PREC_MIXED}[precision]
    p.mode, p.max_iter, p.pin = int(mode), int(max_iter), int(pin)
    p.pixel_border, p.img_w, p.img_h = int(pixel_border), int(img_size[0]), int(img_size[1])
    p.sigma_a, p.sigma_b = float(sigma_a), float(sigma_b)
    p.C_thresh, p.Q_thresh = float(C_thresh), float(Q_thresh)
    p.delta_thresh, p.z_eps = float(delta_thresh), float(z_eps)
    if K is not None:
        Kh = K.detach().cpu().numpy() if isinstance(K, torch.Tensor) else np.asarray(K)
        # gauss_newton_calib.py:70-73: 3x3 matrix or (fx, fy, cx, cy)
        k4 = ([Kh[0, 0], Kh[1, 1], Kh[0, 2], Kh[1, 2]] if Kh.shape == (3, 3) else Kh.flatten()[:4])
        for k in range(4):
            p.K4[k] = float(np.float32(k4[k]))
    return p


def _squeeze_last(x):
    """Accept [..., 1]-trailing or squeezed Cs / valid / Q (gauss_newton.py:84-89)."""
    if x.ndim == 3:
        return x[..., 0]
    return x


def _gn_inputs(Twc, Xs, Cs, ii, jj, idx, valid, Q):
    as_np = _wants_numpy(Twc, Xs, Cs, idx, valid, Q)
    dev = _device(Twc, Xs, Cs, idx, valid, Q)
    Twc_d = _dev(Twc, torch.float32, dev)
    Xs_d = _dev(Xs, torch.float32, dev)
    Cs_d = _squeeze_last(_dev(Cs, torch.float32, dev)).contiguous()
    idx_d = _dev(idx, torch.int32, dev)
    if isinstance(valid, torch.Tensor):
        valid_d = valid.to(dev)
        valid_d = (valid_d if valid_d.dtype in (torch.bool, torch.uint8) else valid_d != 0)
        valid_d = valid_d.to(torch.uint8)
    else:
        valid_d = _dev(np.asarray(valid).astype(bool).view(np.uint8), torch.uint8, dev)
    valid_d = _squeeze_last(valid_d).contiguous()
    Q_d = _squeeze_last(_dev(Q, torch.float32, dev)).contiguous()
    ii_h = (ii.detach().cpu().numpy() if isinstance(ii, torch.Tensor) else np.asarray(ii)).astype(np.int32)
    jj_h = (jj.detach().cpu().numpy() if isinstance(jj, torch.Tensor) else np.asarray(jj)).astype(np.int32)
    return as_np, dev, Twc_d, Xs_d, Cs_d, ii_h, jj_h, idx_d, valid_d, Q_d


def _gauss_newton(params_fn, Twc, Xs, Cs, ii, jj, idx, valid, Q, pin):
    as_np, dev, Twc_d, Xs_d, Cs_d, ii_h, jj_h, idx_d, valid_d, Q_d = _gn_inputs(
        Twc, Xs, Cs, ii, jj, idx, valid, Q)
    num_kf, n = Xs_d.shape[0], Xs_d.shape[1]
    e = len(ii_h)
    # early-outs of gauss_newton.py:66-81 (plan.F == 0 covers them; avoid touching empty tensors)
    if e == 0 or num_kf <= pin:
        return _out(Twc_d.clone(), as_np)
    if idx_d.shape != (e, n) or valid_d.shape != (e, n) or Q_d.shape != (e, n) or Cs_d.shape != (num_kf, n):
        raise M3SError("gauss_newton: inconsistent shapes")
    with torch.cuda.device(dev):
        plan = get_plan(ii_h, jj_h, num_kf, n, pin, dev)
        params = params_fn()
        out = torch.empty_like(Twc_d)
        _lib.check(_lib.load().m3s_gauss_newton(
            plan.handle, C.byref(params), _ptr(Twc_d), _ptr(Xs_d), _ptr(Cs_d), _ptr(idx_d),
            _ptr(valid_d), _ptr(Q_d), _ptr(out), _stream(dev)), "gauss_newton")
    return _out(out, as_np)


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray: float = 0.003,
                      sigma_dist: float = 10.0, C_thresh: float = 0.0, Q_thresh: float = 1.5,
                      max_iter: int = 10, delta_thresh: float = 1e-4, pin: int = 1,
                      use_metal: bool = True, semantics=None):
    """kernels.py:262-322.  ``sigma_dist`` is accepted and unused, as in the reference
    (gauss_newton.py:52)."""
    return _gauss_newton(
        lambda: make_params(_lib.GN_RAYS, max_iter, pin, delta_thresh, C_thresh, Q_thresh, sigma_ray,
                            semantics=semantics),
        Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, pin)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point: float = 0.01,
                        C_thresh: float = 0.0, Q_thresh: float = 1.5, max_iter: int = 10,
                        delta_thresh: float = 1e-4, pin: int = 1, use_metal: bool = True,
                        semantics=None):
    """kernels.py:396-455."""
    return _gauss_newton(
        lambda: make_params(_lib.GN_POINTS, max_iter, pin, delta_thresh, C_thresh, Q_thresh,
                            sigma_point, semantics=semantics),
        Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, pin)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, img_size,
                       pixel_border: int = 0, z_eps: float = 0.0, sigma_pixel: float = 1.0,
                       sigma_depth: float = 0.1, C_thresh: float = 0.0, Q_thresh: float = 1.5,
                       max_iter: int = 10, delta_thresh: float = 1e-4, pin: int = 1,
                       use_metal: bool = True, semantics=None):
    """kernels.py:325-393."""
    return _gauss_newton(
        lambda: make_params(_lib.GN_CALIB, max_iter, pin, delta_thresh, C_thresh, Q_thresh,
                            sigma_pixel, sigma_depth, K, img_size, pixel_border, z_eps,
                            semantics=semantics),
        Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, pin)
