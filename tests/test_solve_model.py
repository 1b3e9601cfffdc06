"""CPU model of the dataflow factorisation of csrc/gn.cu (chol_tiles_kernel): numpy restatement of its
algebra -- block LDL^T with 2x2 pivots giving M_k with M_k A_kk M_k^T = I per 32x32 diagonal block, panels
P = A M^T, the right-hand side carried as one more panel row, back-substitution with M_k^T -- checked against
LU on the oracle's own damped systems.  Pins the maths the kernel implements (gauss_newton.py:254-265 solves
the same system with np.linalg.solve)."""
import numpy as np
import pytest


def elim32(W):
    """tl_elim32_w: M with M W M^T = I from a block LDL^T with 2x2 pivots (no pivoting: W is SPD)."""
    a = W.copy()
    n = np.eye(32)
    M = np.zeros((32, 32))
    for q in range(16):
        j = 2 * q
        al, be, ga = a[j, j], a[j + 1, j], a[j + 1, j + 1]
        det = al * ga - be * be
        assert al > 0 and det > 0
        r = 1.0 / det
        x, y = a[:, j].copy(), a[:, j + 1].copy()
        m1, m2 = (x * ga - y * be) * r, (y * al - x * be) * r      # (x, y) D^-1
        for i in range(j + 2, 32):
            a[i, j + 2:i + 1] -= m1[i] * x[j + 2:i + 1] + m2[i] * y[j + 2:i + 1]
        ra = 1.0 / np.sqrt(al)
        rs = 1.0 / np.sqrt(det * ra * ra)
        nj, nj1 = n[j].copy(), n[j + 1].copy()
        M[j] = ra * nj                                             # F = D^-1/2 = [[ra, 0], [-be ra^2 rs, rs]]
        M[j + 1] = -(be * ra * ra) * rs * nj + rs * nj1
        for i in range(j + 2, 32):
            n[i] -= m1[i] * nj + m2[i] * nj1                       # N = L^-1
    return M


def solve_model(H, g):
    dim = H.shape[0]
    npad = (dim + 31) // 32 * 32
    nblk = npad // 32
    A = np.eye(npad + 1, npad)
    A[:dim, :dim] = H
    A[npad, :] = 0.0
    A[npad, :dim] = -g                                             # the right-hand side as an extra row
    G = np.zeros_like(A)
    Ms = []
    for k in range(nblk):
        s = slice(32 * k, 32 * k + 32)
        Wk = np.tril(A[s, s]) + np.tril(A[s, s], -1).T             # only the lower triangle is kept up to date
        Mk = elim32(Wk)
        assert np.abs(Mk @ Wk @ Mk.T - np.eye(32)).max() < 1e-9
        Ms.append(Mk)
        P = A[32 * k + 32:, s] @ Mk.T                              # panel incl. the rhs row
        G[32 * k + 32:, s] = P
        rows = npad - (32 * k + 32)
        A[32 * k + 32:npad, 32 * k + 32:] -= P[:rows] @ P[:rows].T
        A[npad, 32 * k + 32:] -= P[rows] @ P[:rows].T
    y = G[npad]
    x = np.zeros(npad)
    for b in range(nblk - 1, -1, -1):                              # gn_backsolve_*: x_b = M_b^T (y_b - sum P_rb^T x_r)
        s = slice(32 * b, 32 * b + 32)
        x[s] = Ms[b].T @ (y[s] - G[32 * b + 32:npad, s].T @ x[32 * b + 32:])
    return x[:dim]


@pytest.mark.parametrize("num_kf,loops", [(6, 2), (24, 5), (40, 10)])
@pytest.mark.parametrize("fn", ["gauss_newton_rays", "gauss_newton_points"])
def test_tiled_factorisation_model(oracle, num_kf, loops, fn):
    from mast3r_slam_b200 import synth
    g = synth.make_graph(3, num_kf, loops, 48, 64, num_pts=1500)
    args = (g["Twc"], g["Xs"], g["Cs"], g["ii"], g["jj"], g["idx"], g["valid"], g["Q"])
    _, dbg = getattr(oracle, fn)(*args, max_iter=1, debug=True)
    H = dbg["H"] + 1e-6 * np.eye(dbg["H"].shape[0])               # gauss_newton.py:257 damping
    x = solve_model(H, dbg["g"])
    ref = np.linalg.solve(H, -dbg["g"])
    assert np.abs(x - ref).max() <= 1e-10 * np.abs(ref).max()
    assert np.abs(x - dbg["dx"][0]).max() <= 1e-9 * np.abs(ref).max()
