"""CPU model of the sparse front of csrc/gn.cu (build_sparse_front, sp_level_kernel, sp_tail_kernel,
sp_back_finish_kernel): numpy restatement of the symbolic level selection and of the left-looking 7x7 block
elimination, checked against a dense solve of the same damped system (gauss_newton.py:254-262 calls
np.linalg.solve on it).  Pins the algebra and the list construction the kernels walk."""
import numpy as np
import pytest

from mast3r_slam_b200 import synth

B = 7
MAX_ROWS = 25


def levels_of(adj, min_level=6, max_levels=32):
    """build_sparse_front: greedy independent sets, lowest current degree first, fill edges added."""
    F = len(adj)
    adj = [set(a) for a in adj]
    alive = [True] * F
    struct, enode, level_begin = {}, [], [0]
    n_alive = F
    for _ in range(max_levels):
        cand = sorted((len(adj[v]), v) for v in range(F) if alive[v] and len(adj[v]) <= MAX_ROWS - 1)
        blocked, sel = set(), []
        for _, v in cand:
            if v in blocked or len(sel) >= n_alive - 1:
                continue
            sel.append(v)
            blocked.add(v)
            blocked |= adj[v]
        if len(sel) < min_level:
            break
        for v in sorted(sel):
            struct[v] = sorted(adj[v])
            for a in struct[v]:
                adj[a].discard(v)
                adj[a].update(x for x in struct[v] if x != a)
            adj[v] = set()
            alive[v] = False
            enode.append(v)
            n_alive -= 1
        level_begin.append(len(enode))
    return enode, struct, level_begin, [v for v in range(F) if alive[v]]


def blk(M, a, b):
    return M[B * a:B * a + B, B * b:B * b + B]


def sparse_front_solve(H, rhs, adj, min_level=6):
    enode, struct, level_begin, tail = levels_of(adj, min_level)
    W, Ginv, y = {}, {}, {}
    contrib_of = {v: [] for v in range(len(adj))}
    for j in enode:                                   # level-major order = the order the kernels run in
        At = {a: blk(H, a, j).copy() for a in [j] + struct[j]}
        bt = rhs[B * j:B * j + B].copy()
        for i in contrib_of[j]:                       # earlier-eliminated i with j in struct(i)
            for a in At:
                if a == j or a in struct[i]:
                    At[a] -= W[(a, i)] @ W[(j, i)].T
            bt -= W[(j, i)] @ y[i]
        G = np.linalg.cholesky(At[j])
        Ginv[j] = np.linalg.inv(G)
        for a in struct[j]:
            W[(a, j)] = At[a] @ Ginv[j].T
            contrib_of[a].append(j)
        y[j] = Ginv[j] @ bt
    nR = len(tail)
    Hr, br = np.zeros((B * nR, B * nR)), np.zeros(B * nR)
    for ra, a in enumerate(tail):
        for rb, b in enumerate(tail[:ra + 1]):
            s = blk(H, a, b).copy()
            for i in contrib_of[a]:
                if i in contrib_of[b]:
                    s -= W[(a, i)] @ W[(b, i)].T
            Hr[B * ra:B * ra + B, B * rb:B * rb + B] = s
            Hr[B * rb:B * rb + B, B * ra:B * ra + B] = s.T
        br[B * ra:B * ra + B] = rhs[B * a:B * a + B] - sum((W[(a, i)] @ y[i] for i in contrib_of[a]), np.zeros(B))
    x = np.zeros(len(rhs))
    xr = np.linalg.solve(Hr, br)
    for ra, a in enumerate(tail):
        x[B * a:B * a + B] = xr[B * ra:B * ra + B]
    for l in range(len(level_begin) - 2, -1, -1):
        for j in enode[level_begin[l]:level_begin[l + 1]]:
            t = y[j] - sum((W[(a, j)].T @ x[B * a:B * a + B] for a in struct[j]), np.zeros(B))
            x[B * j:B * j + B] = Ginv[j].T @ t
    return x, dict(levels=len(level_begin) - 1, eliminated=len(enode), tail=nR,
                   max_struct=max((len(s) for s in struct.values()), default=0))


def damped_system(num_kf, loops, seed):
    rng = np.random.default_rng(seed)
    ii, jj = synth.graph_edges(num_kf, loops, rng)
    F = num_kf - 1                                     # keyframe 0 pinned
    H = np.zeros((B * F, B * F))
    adj = [set() for _ in range(F)]
    for a, b in zip(ii[::2], jj[::2]):                 # one undirected pair per two directed rows
        J = rng.standard_normal((12, 2 * B))           # a random edge Jacobian: H_e = J^T J on blocks (a, b)
        He = J.T @ J
        for (u, U) in ((a, 0), (b, 1)):
            for (v, V) in ((a, 0), (b, 1)):
                if u >= 1 and v >= 1:
                    blk(H, u - 1, v - 1)[...] += He[B * U:B * U + B, B * V:B * V + B]
        if a >= 1 and b >= 1:
            adj[a - 1].add(b - 1)
            adj[b - 1].add(a - 1)
    H += 1e-6 * np.eye(B * F)
    return H, rng.standard_normal(B * F), adj


@pytest.mark.parametrize("num_kf,loops,min_level", [(200, 103, 6), (60, 12, 6), (24, 5, 1), (9, 2, 1)])
def test_sparse_front_model_solves_the_damped_system(num_kf, loops, min_level):
    H, rhs, adj = damped_system(num_kf, loops, 3)
    x, info = sparse_front_solve(H, rhs, adj, min_level)
    ref = np.linalg.solve(H, rhs)
    assert np.abs(x - ref).max() <= 1e-9 * np.abs(ref).max()
    assert info["eliminated"] + info["tail"] == num_kf - 1 and info["tail"] >= 1
    assert info["max_struct"] <= MAX_ROWS - 1
    if num_kf == 200:                                   # config-5-like graph: most keyframes leave in the levels
        assert info["levels"] >= 4 and info["tail"] <= 70


def test_sparse_front_model_complete_graph_has_no_levels():
    n = 30
    adj = [set(range(n)) - {v} for v in range(n)]
    enode, _, level_begin, tail = levels_of(adj)
    assert enode == [] and level_begin == [0] and len(tail) == n
