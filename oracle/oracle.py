"""ctypes front-end of the CPU oracle (oracle/m3s_oracle.c).

TEST INFRASTRUCTURE ONLY -- the parity checker and the timed CPU baseline.  Only tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this
module; the product package (mast3r-slam_b200/) never does.

Function signatures mirror the reference's operator API
(/root/reference/src/mlx_mast3r_slam/backends/mpsgraph/kernels.py:107,262,325,396,463) so that
parity tests read like calls into the reference.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libm3s_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc, -ffp-contract=off).  Returns the .so path."""
    src = os.path.join(_HERE, "m3s_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE, "CC=gcc"])
    return _SO


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        _lib = C.CDLL(_SO)
        _lib.m3s_oracle_num_threads.restype = C.c_int
    return _lib


def num_threads() -> int:
    return int(lib().m3s_oracle_num_threads())


def set_threads(n: int) -> None:
    lib().m3s_oracle_set_threads(C.c_int(int(n)))


def _p(a, ty):
    return a.ctypes.data_as(C.POINTER(ty))


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def _u8(a):
    return np.ascontiguousarray(np.asarray(a).astype(bool)).view(np.uint8)


def prep_for_iter_proj(X11, X21, idx_init=None):
    """matching.py:134-175 (fp32).  X11, X21 [B,H,W,3] -> rays [B,H,W,9], pts [B,N,3], p_init [B,N,2]."""
    X11, X21 = _f32(X11), _f32(X21)
    b, h, w, _ = X11.shape
    n = h * w
    rays = np.empty((b, h, w, 9), np.float32)
    pts = np.empty((b, n, 3), np.float32)
    p0 = np.empty((b, n, 2), np.float32)
    ii = None if idx_init is None else _i32(idx_init)
    rc = lib().m3s_oracle_prep_iter_proj(
        _p(X11, C.c_float), _p(X21, C.c_float), None if ii is None else _p(ii, C.c_int32),
        _p(rays, C.c_float), _p(pts, C.c_float), _p(p0, C.c_float), b, h, w)
    assert rc == 0
    return rays, pts, p0


def iter_proj(rays_with_grad, pts3d_norm, p_init, max_iter=10, lambda_init=1e-8,
              convergence_thresh=1e-6, return_iters=False):
    """kernels.py:151-254."""
    rays, pts, p0 = _f32(rays_with_grad), _f32(pts3d_norm), _f32(p_init)
    b, h, w, _ = rays.shape
    n = pts.shape[1]
    p = np.empty((b, n, 2), np.float32)
    valid = np.empty((b, n), np.uint8)
    it = C.c_int(0)
    rc = lib().m3s_oracle_iter_proj(
        _p(rays, C.c_float), _p(pts, C.c_float), _p(p0, C.c_float), _p(p, C.c_float),
        _p(valid, C.c_uint8), b, h, w, n, int(max_iter), C.c_float(lambda_init),
        C.c_float(convergence_thresh), C.byref(it))
    assert rc == 0
    if return_iters:
        return p, valid.astype(bool), it.value
    return p, valid.astype(bool)


def refine_matches(D11, D21, p1, radius=3, dilation_max=0):
    """kernels.py:496-537."""
    D11, D21 = _f32(D11), _f32(D21)
    p1 = np.ascontiguousarray(np.asarray(p1).astype(np.int32))  # int(p1[...]) truncates
    b, h, w, d = D11.shape
    n = p1.shape[1]
    out = np.empty((b, n, 2), np.int32)
    rc = lib().m3s_oracle_refine_matches(
        _p(D11, C.c_float), _p(D21, C.c_float), _p(p1, C.c_int32), _p(out, C.c_int32),
        b, h, w, d, n, int(radius), int(dilation_max))
    assert rc == 0
    return out


def match_tail(X11, X21, p1_int, valid_proj, dist_thresh=0.1):
    """matching.py:438-461 -> (idx [B,N] i32, valid [B,N,1] bool)."""
    X11, X21 = _f32(X11), _f32(X21)
    b, h, w, _ = X11.shape
    n = h * w
    p1 = _i32(p1_int)
    vp = _u8(valid_proj)
    idx = np.empty((b, n), np.int32)
    valid = np.empty((b, n), np.uint8)
    rc = lib().m3s_oracle_match_tail(
        _p(X11, C.c_float), _p(X21, C.c_float), _p(p1, C.c_int32), _p(vp, C.c_uint8),
        _p(idx, C.c_int32), _p(valid, C.c_uint8), b, h, w, C.c_float(dist_thresh))
    assert rc == 0
    return idx, valid.astype(bool)[:, :, None]


def match_iterative_proj(X11, X21, D11, D21, idx_init=None, max_iter=10, lambda_init=1e-8,
                         convergence_thresh=1e-6, radius=3, dilation_max=2, dist_thresh=0.1):
    """matching.py:339-461 end to end on the CPU oracle."""
    b, h, w, _ = np.asarray(X11).shape
    rays, pts, p0 = prep_for_iter_proj(X11, X21, idx_init)
    p, vproj = iter_proj(rays, pts, p0, max_iter, lambda_init, convergence_thresh)
    p1 = p.astype(np.int32)
    if radius > 0:
        p1 = refine_matches(np.asarray(D11).reshape(b, h, w, -1),
                            np.asarray(D21).reshape(b, h * w, -1), p1, radius, dilation_max)
    return match_tail(X11, X21, p1, vproj, dist_thresh)


def match_simple(X11, X21, idx_init=None, dist_thresh=0.1):
    """matching.py:41-90 (the reference's default route, config.py:65 use_simple=True): idx is the
    initial correspondence (identity when None), valid = ||X11[idx] - X21|| < dist_thresh.  Equals the
    match tail with p1 = lin_to_pixel(idx) and valid_proj = all-true.  -> (idx [B,N] i32, valid [B,N,1])."""
    X11, X21 = _f32(X11), _f32(X21)
    b, h, w, _ = X11.shape
    n = h * w
    if idx_init is None:
        idx = np.ascontiguousarray(np.broadcast_to(np.arange(n, dtype=np.int32)[None], (b, n)))
    else:
        idx = _i32(idx_init)
    samp = np.stack([X11.reshape(b, n, 3)[bi, idx[bi]] for bi in range(b)])
    d = (samp - X21.reshape(b, n, 3)) ** 2
    dist = np.sqrt((d[..., 0] + d[..., 1]) + d[..., 2])
    return idx, (dist < np.float32(dist_thresh))[:, :, None]


def _gn(mode, Twc, Xs, Cs, Kmat, ii, jj, idx, valid, Q, img_size, pixel_border, z_eps, sigma_a,
        sigma_b, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug=False):
    Twc, Xs = _f32(Twc), _f32(Xs)
    Cs, Q = np.asarray(Cs), np.asarray(Q)
    valid = np.asarray(valid)
    if Cs.ndim == 3:
        Cs = Cs[..., 0]
    if Q.ndim == 3:
        Q = Q[..., 0]
    if valid.ndim == 3:
        valid = valid[..., 0]
    Cs, Q, valid = _f32(Cs), _f32(Q), _u8(valid)
    ii, jj, idx = _i32(ii), _i32(jj), _i32(idx)
    nkf, n = Xs.shape[0], Xs.shape[1]
    e = len(ii)
    if Kmat is None:
        k4 = np.zeros(4, np.float32)
    else:
        Kmat = np.asarray(Kmat)
        k4 = (np.array([Kmat[0, 0], Kmat[1, 1], Kmat[0, 2], Kmat[1, 2]]) if Kmat.shape == (3, 3)
              else Kmat.flatten()[:4]).astype(np.float32)
    out = np.empty((nkf, 8), np.float32)
    nu = len(np.unique(np.concatenate([ii, jj]))) if e else 0
    dim = 7 * max(nu - pin, 0)
    dx = np.zeros((max(max_iter, 1), max(dim, 1)), np.float64)
    Hg = np.zeros(max(dim * dim + dim, 1), np.float64)
    its = C.c_int(0)
    f = lib().m3s_oracle_gauss_newton(
        int(mode), _p(Twc, C.c_float), _p(Xs, C.c_float), _p(Cs, C.c_float), _p(k4, C.c_float),
        _p(ii, C.c_int32), _p(jj, C.c_int32), _p(idx, C.c_int32), _p(valid, C.c_uint8),
        _p(Q, C.c_float), nkf, C.c_long(n), e, int(img_size[0]), int(img_size[1]),
        int(pixel_border), C.c_double(z_eps), C.c_double(sigma_a), C.c_double(sigma_b),
        C.c_double(C_thresh), C.c_double(Q_thresh), int(max_iter), C.c_double(delta_thresh),
        int(pin), _p(out, C.c_float), _p(dx, C.c_double), _p(Hg, C.c_double), C.byref(its))
    assert f >= 0
    if debug:
        return out, dict(F=f, dx=dx[:, :dim], H=Hg[:dim * dim].reshape(dim, dim),
                         g=Hg[dim * dim:dim * dim + dim], iters=its.value)
    return out


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_ray=0.003,
                      sigma_dist=10.0, C_thresh=0.0, Q_thresh=1.5, max_iter=10, delta_thresh=1e-4,
                      pin=1, debug=False):
    """gauss_newton.py:23-280 (sigma_dist is accepted and unused, as in the reference)."""
    return _gn(0, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, (0, 0), 0, 0.0, sigma_ray,
               0.0, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx_ii2jj, valid_match, Q, sigma_point=0.01,
                        C_thresh=0.0, Q_thresh=1.5, max_iter=10, delta_thresh=1e-4, pin=1,
                        debug=False):
    """gauss_newton_points.py:17-207."""
    return _gn(1, Twc, Xs, Cs, None, ii, jj, idx_ii2jj, valid_match, Q, (0, 0), 0, 0.0,
               sigma_point, 0.0, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, img_size,
                       pixel_border=0, z_eps=0.0, sigma_pixel=1.0, sigma_depth=0.1, C_thresh=0.0,
                       Q_thresh=1.5, max_iter=10, delta_thresh=1e-4, pin=1, debug=False):
    """gauss_newton_calib.py:17-274."""
    return _gn(2, Twc, Xs, Cs, K, ii, jj, idx_ii2jj, valid_match, Q, img_size, pixel_border, z_eps,
               sigma_pixel, sigma_depth, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug)


# ---------------------------------------------------------------------------------------------
# numpy restatements of the MLX-only host steps (cannot be imported from the reference: they need
# `mlx`).  Used to cross-check the C versions above at small sizes.
# ---------------------------------------------------------------------------------------------
def prep_for_iter_proj_numpy(X11, X21, idx_init=None):
    """matching.py:134-175 + image.py:9-34 restated with numpy fp32 ops."""
    X11 = np.asarray(X11, np.float32)
    X21 = np.asarray(X21, np.float32)
    b, h, w, _ = X11.shape

    def normalize(X):
        sq = X * X
        s = (sq[..., 0] + sq[..., 1]) + sq[..., 2]
        return X / np.sqrt(s + np.float32(1e-10))[..., None]

    rays = normalize(X11)
    gx = np.zeros_like(rays)
    gy = np.zeros_like(rays)
    gx[:, :, 1:-1] = (rays[:, :, 2:] - rays[:, :, :-2]) / np.float32(2.0)
    gy[:, 1:-1] = (rays[:, 2:] - rays[:, :-2]) / np.float32(2.0)
    rwg = np.concatenate([rays, gx, gy], axis=-1)
    pts = normalize(X21.reshape(b, -1, 3))
    if idx_init is None:
        idx_init = np.broadcast_to(np.arange(h * w)[None], (b, h * w))
    idx_init = np.asarray(idx_init)
    p0 = np.stack([idx_init % w, idx_init // w], axis=-1).astype(np.float32)
    return rwg, pts, p0


def match_tail_numpy(X11, X21, p1_int, valid_proj, dist_thresh=0.1):
    """matching.py:438-461 restated with numpy."""
    X11 = np.asarray(X11, np.float32)
    X21 = np.asarray(X21, np.float32)
    b, h, w, _ = X11.shape
    p1_int = np.asarray(p1_int, np.int32)
    y = np.clip(p1_int[..., 1], 0, h - 1)
    x = np.clip(p1_int[..., 0], 0, w - 1)
    lin = y * w + x
    X11f = X11.reshape(b, h * w, 3)
    samp = np.stack([X11f[bi, lin[bi]] for bi in range(b)])
    d = (samp - X21.reshape(b, h * w, 3)) ** 2
    dist = np.sqrt((d[..., 0] + d[..., 1]) + d[..., 2])
    valid = np.asarray(valid_proj, bool) & (dist < np.float32(dist_thresh))
    idx = p1_int[..., 0] + w * p1_int[..., 1]
    return idx.astype(np.int32), valid[:, :, None]


def constrain_points_to_ray_numpy(img_size, Xs, K):
    """geometry.py:273-302 + backproject :229-246 restated with numpy fp32 ops (MLX fp32 in the
    reference; cannot be imported here).  img_size = (H, W); K 3x3 or (fx, fy, cx, cy)."""
    Xs = np.asarray(Xs, np.float32)
    K = np.asarray(K, np.float32)
    fx, fy, cx, cy = (K[0, 0], K[1, 1], K[0, 2], K[1, 2]) if K.shape == (3, 3) else K
    h, w = img_size
    v, u = np.meshgrid(np.arange(h, dtype=np.float32), np.arange(w, dtype=np.float32), indexing="ij")
    u = np.broadcast_to(u.reshape(1, -1, 1), Xs[..., 2:3].shape)
    v = np.broadcast_to(v.reshape(1, -1, 1), Xs[..., 2:3].shape)
    z = Xs[..., 2:3]
    x = (u - np.float32(cx)) / np.float32(fx) * z
    y = (v - np.float32(cy)) / np.float32(fy) * z
    return np.concatenate([x, y, z], axis=-1)


def add_factors_numpy(ii, jj, match_out, Q_conf, min_match_frac, is_reloc=False):
    """FactorGraph.add_factors (global_opt.py:86-136) restated with numpy fp32: returns None when the
    reference returns False, else the rows it appends
    (ii, jj, idx_i2j, idx_j2i, valid_match_j, valid_match_i, Qj, Qi)."""
    idx_i2j, idx_j2i, vmj, vmi, Qii, Qjj, Qji, Qij = [np.asarray(a) for a in match_out]
    b = idx_i2j.shape[0]
    bi = np.broadcast_to(np.arange(b)[:, None], idx_i2j.shape)
    Qj = np.sqrt(Qii.astype(np.float32)[bi, idx_i2j] * Qji.astype(np.float32))
    Qi = np.sqrt(Qjj.astype(np.float32)[bi, idx_j2i] * Qij.astype(np.float32))
    valid_j = vmj.astype(bool) & (Qj > Q_conf)
    valid_i = vmi.astype(bool) & (Qi > Q_conf)
    nj = valid_j.shape[1] * valid_j.shape[2]
    ni = valid_i.shape[1] * valid_i.shape[2]
    fj = valid_j.astype(np.float32).sum(axis=(1, 2)) / nj
    fi = valid_i.astype(np.float32).sum(axis=(1, 2)) / ni
    ii_t, jj_t = np.asarray(ii, np.int32), np.asarray(jj, np.int32)
    invalid = np.minimum(fj, fi) < min_match_frac
    invalid = (~(ii_t == (jj_t - 1))) & invalid
    if invalid.any() and is_reloc:
        return None
    keep = ~invalid
    if keep.sum() == 0:
        return None
    return (ii_t[keep], jj_t[keep], idx_i2j[keep], idx_j2i[keep], vmj[keep], vmi[keep], Qj[keep], Qi[keep])


# ---------------------------------------------------------------------------------------------
# Tracker single-pose solves.  numpy restatement of tracker.py:216-406,
# geometry.py:46-227, liegroups/{so3,sim3}.py and optimizer.py:11-62.  The reference evaluates
# these in MLX (not importable here: parity UNPINNED for this row -- restated, never executed);
# `dtype` selects the arithmetic (float64 = the exact-math statement the CUDA path is checked
# against, float32 = the reference's own precision, used to size the tolerance).
# ---------------------------------------------------------------------------------------------
def _cross(a, b):  # liegroups/so3.py:9-25
    return np.stack([a[..., 1] * b[..., 2] - a[..., 2] * b[..., 1],
                     a[..., 2] * b[..., 0] - a[..., 0] * b[..., 2],
                     a[..., 0] * b[..., 1] - a[..., 1] * b[..., 0]], axis=-1)


def _so3_act(q, p):  # so3.py:147-167
    t = 2.0 * _cross(q[..., :3], p)
    return p + q[..., 3:4] * t + _cross(q[..., :3], t)


def _so3_mul(q1, q2):  # so3.py:125-145
    x1, y1, z1, w1 = q1
    x2, y2, z2, w2 = q2
    return np.array([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2],
                    dtype=q1.dtype)


def _sim3_mul(a, b):  # sim3.py:206-220
    t = a[:3] + a[7] * _so3_act(a[3:7], b[:3])
    return np.concatenate([t, _so3_mul(a[3:7], b[3:7]), [a[7] * b[7]]]).astype(a.dtype)


def _sim3_inv(a):  # sim3.py:195-204
    qi = np.concatenate([-a[3:6], a[6:7]])
    si = 1.0 / (a[7] + a.dtype.type(1e-10))
    return np.concatenate([-si * _so3_act(qi, a[:3]), qi, [si]]).astype(a.dtype)


def _sim3_exp(tau):  # sim3.py:106-154 + so3.py:65-96
    dt = tau.dtype.type
    v, w, sigma = tau[:3], tau[3:6], tau[6]
    th2 = np.sum(w * w)
    th = np.sqrt(th2 + dt(1e-10))
    small = th2 < 1e-8
    with np.errstate(all="ignore"):
        A = dt(1.0) - th2 / dt(6.0) if small else np.sin(th) / th
        B = dt(0.5) - th2 / dt(24.0) if small else (dt(1.0) - np.cos(th)) / th2
        Cc = dt(1.0 / 6.0) - th2 / dt(120.0) if small else (dt(1.0) - A) / th2
        wv = _cross(w, v)
        t = v + B * wv + Cc * _cross(w, wv)
        half = dt(0.5) * th
        sh = dt(0.5) - th2 / dt(48.0) if small else np.sin(half) / th
        ch = dt(1.0) - th2 / dt(8.0) if small else np.cos(half)
    return np.concatenate([t, sh * w, [ch], [np.exp(sigma)]]).astype(tau.dtype)


def _track_solve(sqrt_info, r, J, huber):  # tracker.py:216-253
    wr = np.abs(sqrt_info * r)
    with np.errstate(all="ignore"):
        hw = np.where(wr < huber, np.ones_like(wr), huber / wr)
    rs = sqrt_info * np.sqrt(hw)
    m = J.shape[-1]
    A = (rs[..., None] * J).reshape(-1, m)
    b = (rs * r).reshape(-1, 1)
    H = A.T @ A
    g = -(A.T @ b)
    cost = 0.5 * float((b.T @ b).item())
    tau = np.linalg.solve(H + 1e-6 * np.eye(m, dtype=H.dtype), g[:, 0])  # linalg.py:17-50
    return tau.astype(J.dtype), cost


def _converged(rel_thr, dn_thr, old, new, tau):  # optimizer.py:11-46
    with np.errstate(all="ignore"):
        rel = abs(np.float64(old - new) / np.float64(old + 1e-10))
    dn = float(np.sqrt(np.sum(tau.astype(np.float64) ** 2)))
    return bool(rel < rel_thr or dn < dn_thr)


def _act_jac(T, X):  # geometry.py:95-137
    pW = T[7] * _so3_act(T[3:7][None], X) + T[:3]
    n = pW.shape[0]
    J = np.zeros((n, 3, 7), dtype=pW.dtype)
    J[:, 0, 0] = J[:, 1, 1] = J[:, 2, 2] = 1.0
    x, y, z = pW[:, 0], pW[:, 1], pW[:, 2]
    J[:, 0, 4], J[:, 0, 5] = z, -y          # -skew(pW)
    J[:, 1, 3], J[:, 1, 5] = -z, x
    J[:, 2, 3], J[:, 2, 4] = y, -x
    J[:, :, 6] = pW
    return pW, J


def _ray_dist(X, jac):  # geometry.py:46-92
    d = np.sqrt(np.sum(X * X, axis=-1, keepdims=True) + X.dtype.type(1e-10))
    di = 1.0 / d
    r = di * X
    rd = np.concatenate([r, d], axis=-1)
    if not jac:
        return rd
    eye = np.eye(3, dtype=X.dtype)[None]
    dr = di[:, :, None] * (eye - (di ** 2)[:, :, None] * (X[:, :, None] @ X[:, None, :]))
    return rd, np.concatenate([dr, r[:, None, :]], axis=1)


DEFAULT_TRACKING = {"C_conf": 0.0, "Q_conf": 1.5, "rel_error": 1e-3, "delta_norm": 1e-3,
                    "max_iters": 10, "huber": 1.345, "sigma_ray": 0.003, "sigma_dist": 10.0,
                    "sigma_pixel": 1.0, "sigma_depth": 10.0, "pixel_border": 0, "depth_eps": 0.0}


def opt_pose_ray_dist_sim3_numpy(Xf, Xk, T_WCf, T_WCk, Qk, valid, cfg=None, dtype=np.float64):
    """tracker.py:255-313.  Returns (T_WCf, T_CkCf, info)."""
    cfg = {**DEFAULT_TRACKING, **(cfg or {})}
    Xf, Xk = np.asarray(Xf, dtype).reshape(-1, 3), np.asarray(Xk, dtype).reshape(-1, 3)
    Qk = np.asarray(Qk, dtype).reshape(-1, 1)
    vf = np.asarray(valid).reshape(-1, 1).astype(dtype)
    Tf, Tk = np.asarray(T_WCf, dtype).reshape(8), np.asarray(T_WCk, dtype).reshape(8)
    si_ray = dtype(1.0 / cfg["sigma_ray"]) * vf * np.sqrt(Qk)
    si_dist = dtype(1.0 / cfg["sigma_dist"]) * vf * np.sqrt(Qk)
    sqrt_info = np.concatenate([np.broadcast_to(si_ray, (len(Qk), 3)), si_dist], axis=-1)
    T = _sim3_mul(_sim3_inv(Tk), Tf)
    rd_k = _ray_dist(Xk, False)
    old, it, cost = float("inf"), 0, 0.0
    for _ in range(cfg["max_iters"]):
        pW, dX = _act_jac(T, Xf)
        rd_f, drd = _ray_dist(pW, True)
        r = rd_k - rd_f
        J = -(drd @ dX)
        tau, cost = _track_solve(sqrt_info, r, J, dtype(cfg["huber"]))
        T = _sim3_mul(T, _sim3_exp(tau))
        it += 1
        if _converged(cfg["rel_error"], cfg["delta_norm"], old, cost, tau):
            break
        old = cost
    return _sim3_mul(Tk, T), T, {"iterations": it, "cost": cost,
                                 "delta_norm": float(np.linalg.norm(tau)) if it else 0.0}


def opt_pose_calib_sim3_numpy(Xf, Xk, T_WCf, T_WCk, Qk, valid, K, img_size, cfg=None, dtype=np.float64):
    """tracker.py:315-406 with meas_k / valid_meas_k from _get_points_poses (:195-212) and
    project_calib (geometry.py:155-227).  img_size = (h, w); K = 3x3 or (fx,fy,cx,cy)."""
    cfg = {**DEFAULT_TRACKING, **(cfg or {})}
    K = np.asarray(K, dtype)
    fx, fy, cx, cy = (K[0, 0], K[1, 1], K[0, 2], K[1, 2]) if K.shape == (3, 3) else K
    h, w = int(img_size[0]), int(img_size[1])
    Xf, Xk = np.asarray(Xf, dtype).reshape(-1, 3), np.asarray(Xk, dtype).reshape(-1, 3)
    n = len(Xk)
    Qk = np.asarray(Qk, dtype).reshape(-1, 1)
    vf = np.asarray(valid).reshape(-1, 1).astype(dtype)
    Tf, Tk = np.asarray(T_WCf, dtype).reshape(8), np.asarray(T_WCk, dtype).reshape(8)
    border, z_eps = cfg["pixel_border"], cfg["depth_eps"]
    lin = np.arange(n)
    with np.errstate(all="ignore"):
        meas = np.stack([(lin % w).astype(dtype), (lin // w).astype(dtype),
                         np.log(Xk[:, 2] + dtype(1e-10))], axis=-1)
    vmeas = Xk[:, 2:3] > z_eps
    meas = np.where(vmeas, meas, 0)
    si_px = dtype(1.0 / cfg["sigma_pixel"]) * vf * np.sqrt(Qk)
    si_z = dtype(1.0 / cfg["sigma_depth"]) * vf * np.sqrt(Qk)
    sqrt_info = np.concatenate([np.broadcast_to(si_px, (n, 2)), si_z], axis=-1)
    T = _sim3_mul(_sim3_inv(Tk), Tf)
    old, it, cost = float("inf"), 0, 0.0
    for _ in range(cfg["max_iters"]):
        P, dX = _act_jac(T, Xf)
        x, y, z = P[:, 0], P[:, 1], P[:, 2]
        with np.errstate(all="ignore"):
            zden = z + dtype(1e-10)
            u, v = (fx * x + cx * z) / zden, (fy * y + cy * z) / zden
            vz = z > z_eps
            vproj = (u > border) & (u < w - 1 - border) & (v > border) & (v < h - 1 - border) & vz
            logz = np.where(vz, np.log(z + dtype(1e-10)), 0)
            zi = 1.0 / zden
        dpz = np.zeros((n, 3, 3), dtype=dtype)
        dpz[:, 0, 0], dpz[:, 0, 2] = fx * zi, -fx * x * zi * zi
        dpz[:, 1, 1], dpz[:, 1, 2] = fy * zi, -fy * y * zi * zi
        dpz[:, 2, 2] = zi
        v2 = (vproj[:, None] & vmeas)
        si2 = np.where(v2, sqrt_info, 0)
        r = np.where(v2, meas - np.stack([u, v, logz], axis=-1), 0)   # masked rows carry weight 0
        J = np.where(v2[..., None], -(dpz @ dX), 0)
        tau, cost = _track_solve(si2, r, J, dtype(cfg["huber"]))
        T = _sim3_mul(T, _sim3_exp(tau))
        it += 1
        if _converged(cfg["rel_error"], cfg["delta_norm"], old, cost, tau):
            break
        old = cost
    return _sim3_mul(Tk, T), T, {"iterations": it, "cost": cost,
                                 "delta_norm": float(np.linalg.norm(tau)) if it else 0.0}


def track_gate_numpy(Cf, Ck, Qff, Qkf, idx_f2k, valid_match_k, C_conf=0.0, Q_conf=1.5):
    """tracker.py:86-113 in float32."""
    idx = np.asarray(idx_f2k).reshape(-1)
    Cf, Ck = _f32(Cf).reshape(-1), _f32(Ck).reshape(-1)
    Qff, Qkf = _f32(Qff).reshape(-1), _f32(Qkf).reshape(-1)
    vm = np.asarray(valid_match_k).reshape(-1).astype(bool)
    Qk = np.sqrt(Qff[idx] * Qkf)
    v_q = Qk > np.float32(Q_conf)
    v_opt = vm & (Cf[idx] > np.float32(C_conf)) & (Ck > np.float32(C_conf)) & v_q
    v_kf = vm & v_q
    return Qk[:, None], v_opt[:, None], v_kf[:, None], (int(v_opt.sum()), int(v_kf.sum()))
