"""Oracle of the Metal-semantics mode.  TEST INFRASTRUCTURE ONLY -- only tests/,
__graft_entry__.smoke() and the golden generator may import this.

The per-thread arithmetic is the reference's own: oracle/_ref/libm3s_metal_ref.so is built (oracle/
Makefile, target `metal_ref`) from the UNMODIFIED shaders under
/root/reference/src/mlx_mast3r_slam/backends/mpsgraph/shaders compiled as C++ through
oracle/metal_shim/metal_stdlib.  What is restated here, in numpy, is the HOST side of the reference's
Metal runners, line by line:

    iter_proj        metal_runner.py:118-196          one dispatch, thread per point
    refine_matches   refine_metal_runner.py:100-164   multiscale kernel, in place, dilation_max..1
    gauss_newton_*   gn_metal_runner.py:96-341, gn_points_metal_runner.py:96-347,
                     gn_calib_metal_runner.py:95-385  per iteration: Jacobian kernel -> float32
                     per-keyframe sums accumulated into float64 H, g -> np.linalg.solve ->
                     pose_update_kernel (float32, dt = tau) on the float32 pose buffer
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(HERE, "_ref", "libm3s_metal_ref.so")
REF_SHADERS = "/root/reference/src/mlx_mast3r_slam/backends/mpsgraph/shaders"
_lib = None


def available() -> bool:
    return os.path.exists(SO) or os.path.isdir(REF_SHADERS)


def build() -> str | None:
    """(Re)build from the reference's sources when they are present (this container); on the GPU box
    the prebuilt library that travelled with the snapshot is used."""
    if os.path.isdir(REF_SHADERS):
        subprocess.check_call(["make", "-s", "-C", HERE, "metal_ref"])
    return SO if os.path.exists(SO) else None


def lib():
    global _lib
    if _lib is None:
        if build() is None:
            raise RuntimeError("oracle/_ref/libm3s_metal_ref.so is missing and /root/reference is not here")
        _lib = C.CDLL(SO)
    return _lib


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def iter_proj(rays, pts, p_init, max_iter=10, lambda_init=1e-8, convergence_thresh=1e-6):
    """metal_runner.py:118-196 -> (p [B,N,2] f32, valid [B,N] bool)"""
    rays, pts, p_init = _f32(rays), _f32(pts), _f32(p_init)
    b, h, w, _ = rays.shape
    n = pts.shape[1]
    p = np.empty((b, n, 2), np.float32)
    valid = np.zeros((b, n), np.uint8)
    lib().mref_iter_proj(_p(rays), _p(pts), _p(p_init), _p(p), _p(valid), b, h, w, n, int(max_iter),
                         C.c_float(lambda_init), C.c_float(convergence_thresh))
    return p, valid.astype(bool)


def refine_matches(D11, D21, p1, radius=3, dilation_max=1):
    """refine_metal_runner.py:100-164 -> p1 [B,N,2] i32"""
    D11, D21 = _f32(D11), _f32(D21)
    b, h, w, d = D11.shape
    n = D21.shape[1]
    out = _i32(p1).copy()
    lib().mref_refine_matches(_p(D11), _p(D21), _p(out), b, h, w, d, n, int(radius), int(dilation_max))
    return out


_TRIU = np.array([(i, j) for i in range(7) for j in range(i, 7)])


def _gauss_newton(mode, Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_a, sigma_b, K4, img_size, pixel_border,
                  z_eps, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug=None):
    Twc = _f32(Twc)
    num_kf, num_edges = Twc.shape[0], len(ii)
    if num_edges == 0 or num_kf <= pin:
        return Twc.copy()
    ii, jj = _i32(ii), _i32(jj)
    unique_kf = np.unique(np.concatenate([ii, jj]))
    if len(unique_kf) <= pin:
        return Twc.copy()
    kf_to_local = {int(kf): i - pin for i, kf in enumerate(unique_kf)}
    num_free = len(unique_kf) - pin
    valid = np.asarray(valid)
    Q = np.asarray(Q)
    Cs = np.asarray(Cs)
    if valid.ndim == 3:
        valid = valid[..., 0]
    if Q.ndim == 3:
        Q = Q[..., 0]
    if Cs.ndim == 3:
        Cs = Cs[..., 0]
    Xs, Cs, Q, idx = _f32(Xs), _f32(Cs), _f32(Q), _i32(idx)
    valid = np.ascontiguousarray(valid.astype(np.uint8))
    num_pts = Xs.shape[1]
    total = num_edges * num_pts
    K4 = _f32(K4 if K4 is not None else np.zeros(4))
    w_img, h_img = (int(img_size[0]), int(img_size[1])) if img_size is not None else (0, 0)
    Twc_buf = Twc.copy()                           # the float32 pose buffer that lives on the GPU
    free_kf = _i32(unique_kf[pin:])
    edge_of = np.arange(total) // num_pts
    i_local_all = np.array([kf_to_local.get(int(x), -pin - 1) for x in ii[edge_of]])
    j_local_all = np.array([kf_to_local.get(int(x), -pin - 1) for x in jj[edge_of]])
    L = lib()
    for _ in range(max_iter):
        JtJ_i = np.zeros((total, 28), np.float32)
        Jtr_i = np.zeros((total, 7), np.float32)
        JtJ_j = np.zeros((total, 28), np.float32)
        Jtr_j = np.zeros((total, 7), np.float32)
        JtJ_ij = np.zeros((total, 49), np.float32)
        vout = np.zeros(total, np.uint8)
        L.mref_gn_jacobian(mode, _p(Twc_buf), _p(Xs), _p(Cs), _p(ii), _p(jj), _p(idx), _p(valid), _p(Q), _p(K4),
                           _p(JtJ_i), _p(Jtr_i), _p(JtJ_j), _p(Jtr_j), _p(JtJ_ij), _p(vout), num_kf, num_pts,
                           num_edges, C.c_float(np.float32(1.0 / sigma_a)),
                           C.c_float(np.float32(1.0 / sigma_b) if sigma_b else 0.0),
                           C.c_float(C_thresh), C.c_float(Q_thresh), w_img, h_img, int(pixel_border),
                           C.c_float(z_eps))
        dim = 7 * num_free
        H = np.zeros((dim, dim), np.float64)
        g = np.zeros(dim, np.float64)
        vi = np.where(vout.astype(bool))[0]
        if len(vi) > 0:                            # gn_metal_runner.py:241-296
            il, jl = i_local_all[vi], j_local_all[vi]
            A_i, A_j, A_ij, b_i, b_j = JtJ_i[vi], JtJ_j[vi], JtJ_ij[vi], Jtr_i[vi], Jtr_j[vi]
            for kf in range(num_free):
                for loc, A, b in ((il, A_i, b_i), (jl, A_j, b_j)):
                    m = loc == kf
                    if np.any(m):
                        s = A[m].sum(axis=0)       # float32 sums, as in the reference
                        for k, (r, c) in enumerate(_TRIU):
                            H[kf * 7 + r, kf * 7 + c] += s[k]
                            if r != c:
                                H[kf * 7 + c, kf * 7 + r] += s[k]
                        g[kf * 7:(kf + 1) * 7] += b[m].sum(axis=0)
            for a in range(num_free):
                for c in range(a + 1, num_free):
                    m_ac = (il == a) & (jl == c)
                    m_ca = (il == c) & (jl == a)
                    if np.any(m_ac):
                        s = A_ij[m_ac].sum(axis=0).reshape(7, 7)
                        H[a * 7:(a + 1) * 7, c * 7:(c + 1) * 7] += s
                        H[c * 7:(c + 1) * 7, a * 7:(a + 1) * 7] += s.T
                    if np.any(m_ca):
                        s = A_ij[m_ca].sum(axis=0).reshape(7, 7)
                        H[c * 7:(c + 1) * 7, a * 7:(a + 1) * 7] += s
                        H[a * 7:(a + 1) * 7, c * 7:(c + 1) * 7] += s.T
        H += np.eye(dim) * 1e-6
        if debug is not None and "H" not in debug:
            debug["H"], debug["g"] = H.copy(), g.copy()
        try:
            dx = np.linalg.solve(H, -g)
        except np.linalg.LinAlgError:
            break
        if np.linalg.norm(dx) < delta_thresh:
            break
        dx32 = _f32(dx)
        L.mref_pose_update(_p(Twc_buf), _p(dx32), _p(free_kf), num_free)
    return Twc_buf


def gauss_newton_rays(Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_ray=0.003, sigma_dist=10.0, C_thresh=0.0,
                      Q_thresh=1.5, max_iter=10, delta_thresh=1e-4, pin=1, debug=None):
    """gn_metal_runner.py:81-341"""
    return _gauss_newton(0, Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_ray, None, None, None, 0, 0.0, C_thresh,
                         Q_thresh, max_iter, delta_thresh, pin, debug)


def gauss_newton_points(Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_point=0.01, C_thresh=0.0, Q_thresh=1.5,
                        max_iter=10, delta_thresh=1e-4, pin=1, debug=None):
    """gn_points_metal_runner.py:91-347"""
    return _gauss_newton(1, Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_point, None, None, None, 0, 0.0, C_thresh,
                         Q_thresh, max_iter, delta_thresh, pin, debug)


def gauss_newton_calib(Twc, Xs, Cs, K, ii, jj, idx, valid, Q, img_size, pixel_border=0, z_eps=0.0,
                       sigma_pixel=1.0, sigma_depth=0.1, C_thresh=0.0, Q_thresh=1.5, max_iter=10,
                       delta_thresh=1e-4, pin=1, debug=None):
    """gn_calib_metal_runner.py:94-385"""
    K = np.asarray(K)
    K4 = np.array([K[0, 0], K[1, 1], K[0, 2], K[1, 2]], np.float32) if K.shape == (3, 3) else _f32(K).ravel()[:4]
    return _gauss_newton(2, Twc, Xs, Cs, ii, jj, idx, valid, Q, sigma_pixel, sigma_depth, K4, img_size,
                         pixel_border, z_eps, C_thresh, Q_thresh, max_iter, delta_thresh, pin, debug)
