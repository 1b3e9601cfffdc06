"""Seeded synthetic inputs for the hot path (numpy only; no reference code involved).

Shapes and value ranges follow SURVEY.md s8(d) / BASELINE.json configs:
  * make_pair      -- one frame pair for matching (configs 1, 2): an analytic surface seen by view 1,
                      and the same surface re-sampled along a smooth few-pixel flow for view 2.
  * make_graph     -- a physically consistent keyframe graph for the Gauss-Newton solvers
                      (configs 3, 4, 5): window-2 chain plus loop closures, both directions.
  * reference_like_* -- re-creations of the value distributions used by the reference's own
                      benchmark generators (benchmark_all_kernels.py:16-42, 67-76, 385-395):
                      incoherent worst cases.
"""
from __future__ import annotations

import numpy as np


def intrinsics(h: int, w: int):
    """Pinhole intrinsics scaled from the 512x384 case K=(420, 420, 255.5, 191.5)."""
    f = 420.0 * w / 512.0
    return f, f, (w - 1) / 2.0, (h - 1) / 2.0


def _surface(u, v, h, w):
    """X(u, v) = z(u, v) * K^-1 [u, v, 1] with z = 2 + 0.5 sin(2 pi u/W) cos(2 pi v/H) + 0.3 u/W."""
    fx, fy, cx, cy = intrinsics(h, w)
    z = 2.0 + 0.5 * np.sin(2 * np.pi * u / w) * np.cos(2 * np.pi * v / h) + 0.3 * (u / w)
    return np.stack([(u - cx) / fx * z, (v - cy) / fy * z, z], axis=-1)


def _smooth_field(rng, h, w, d, cell=6):
    """Smooth d-channel random field: coarse Gaussian noise, bilinearly upsampled.  Returned as a
    callable sampler at continuous (u, v) so that view-2 descriptors are the SAME field evaluated
    at the true correspondence."""
    gh, gw = h // cell + 3, w // cell + 3
    coarse = rng.standard_normal((gh, gw, d)).astype(np.float64)

    def sample(u, v):
        gu = np.clip(u / cell + 1.0, 0, gw - 1.001)
        gv = np.clip(v / cell + 1.0, 0, gh - 1.001)
        x0 = np.floor(gu).astype(np.int64)
        y0 = np.floor(gv).astype(np.int64)
        fx = (gu - x0)[..., None]
        fy = (gv - y0)[..., None]
        return ((1 - fx) * (1 - fy) * coarse[y0, x0] + (1 - fx) * fy * coarse[y0 + 1, x0]
                + fx * (1 - fy) * coarse[y0, x0 + 1] + fx * fy * coarse[y0 + 1, x0 + 1])

    return sample


def make_pair(seed: int = 0, h: int = 384, w: int = 512, d: int = 24, desc: str = "dyadic",
              noise: float = 1e-3):
    """One matching problem.  Returns dict(X11, X21 [H,W,3] f32; D11, D21 [H,W,d] f32;
    flow_uv [H,W,2] the true sub-pixel correspondence of every view-2 pixel in view 1).

    desc = "dyadic": descriptors rounded to k/64, |.|<=2 -> every d-term dot product is exact in
    fp32 under ANY summation order (unconditional bit-exactness of the argmax);
    desc = "unit": un-quantised, L2-normalised (near-ties possible, tests enumerate them)."""
    rng = np.random.default_rng(1000 + seed)
    vv, uu = np.meshgrid(np.arange(h, dtype=np.float64), np.arange(w, dtype=np.float64),
                         indexing="ij")
    X11 = _surface(uu, vv, h, w)
    # smooth flow: similarity about the image centre + a few-pixel shift, per-seed
    s = 1.0 + 0.01 * rng.uniform(-1, 1)
    th = 0.01 * rng.uniform(-1, 1)
    t = rng.uniform(-3.0, 3.0, size=2)
    cx, cy = (w - 1) / 2.0, (h - 1) / 2.0
    du, dv = uu - cx, vv - cy
    u2 = cx + s * (np.cos(th) * du - np.sin(th) * dv) + t[0]
    v2 = cy + s * (np.sin(th) * du + np.cos(th) * dv) + t[1]
    X21 = _surface(u2, v2, h, w) + noise * rng.standard_normal((h, w, 3))
    field = _smooth_field(rng, h, w, d)
    D11 = field(uu, vv) + 0.15 * rng.standard_normal((h, w, d))
    D21 = field(u2, v2) + 0.15 * rng.standard_normal((h, w, d))
    if desc == "dyadic":
        D11 = np.clip(np.round(D11 * 64.0) / 64.0, -2.0, 2.0)
        D21 = np.clip(np.round(D21 * 64.0) / 64.0, -2.0, 2.0)
    elif desc == "unit":
        D11 = D11 / np.linalg.norm(D11, axis=-1, keepdims=True)
        D21 = D21 / np.linalg.norm(D21, axis=-1, keepdims=True)
    else:
        raise ValueError(desc)
    f32 = np.float32
    return dict(X11=X11.astype(f32), X21=X21.astype(f32), D11=D11.astype(f32), D21=D21.astype(f32),
                flow_uv=np.stack([u2, v2], -1).astype(f32))


def make_pairs(seeds, h=384, w=512, d=24, desc="dyadic"):
    """Batch of pairs stacked on a leading B axis (config 2: seeds 0..63)."""
    ps = [make_pair(s, h, w, d, desc) for s in seeds]
    return {k: np.stack([p[k] for p in ps]) for k in ("X11", "X21", "D11", "D21")}


# ---------------------------------------------------------------------------------------------
# Sim3 helpers (generator-side only)
# ---------------------------------------------------------------------------------------------
def _quat_from_rotvec(w):
    th = np.linalg.norm(w)
    if th < 1e-12:
        return np.array([0, 0, 0, 1.0])
    a = w / th
    return np.concatenate([a * np.sin(th / 2), [np.cos(th / 2)]])


def _quat_mul(a, b):
    x1, y1, z1, w1 = a
    x2, y2, z2, w2 = b
    return np.array([w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2, w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2,
                     w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2, w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2])


def _rotmat(q):
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def graph_edges(num_kf: int, num_loops: int, rng):
    """Undirected pairs: window-2 chain (k,k+1),(k,k+2) plus `num_loops` random loop closures,
    returned as directed rows in both directions, pair by pair ((i,j) then (j,i) adjacent)."""
    pairs = []
    for k in range(num_kf - 1):
        pairs.append((k, k + 1))
        if k + 2 < num_kf:
            pairs.append((k, k + 2))
    seen = set(pairs)
    tries = 0
    while num_loops > 0 and tries < 100000:
        tries += 1
        a, b = sorted(rng.integers(0, num_kf, size=2).tolist())
        if b - a < 3 or (a, b) in seen:
            continue
        seen.add((a, b))
        pairs.append((a, b))
        num_loops -= 1
    ii, jj = [], []
    for a, b in pairs:
        ii += [a, b]
        jj += [b, a]
    return np.array(ii, np.int32), np.array(jj, np.int32)


def make_graph(seed: int = 0, num_kf: int = 10, num_loops: int = 3, h: int = 384, w: int = 512,
               num_pts: int | None = None, noise: float = 1e-3, sigma_t: float = 0.02,
               sigma_w: float = 0.01, sigma_s: float = 0.01, valid_frac: float = 0.85):
    """Physically consistent keyframe graph.

    World points P = the analytic surface on the HxW grid (or its first `num_pts` points of a
    strided sub-sampling).  Keyframe k has ground-truth pose T_k (small motion around identity) and
    observes the world points in a circularly shifted order: Xs[k][n] = T_k^-1 P[(n + sh_k) % N]
    (+ noise), so idx_ii2jj[e][n] = (n + sh_j - sh_i) % N is a coherent (contiguous-with-offset)
    gather, like a real flow field.  Poses handed to the solver are the ground truth perturbed by
    (sigma_t, sigma_w, sigma_s); KF 0 is exact (it is the pinned one).

    Returns dict(Twc [K,8] f32 perturbed, Twc_gt, Xs [K,N,3] f32, Cs [K,N] f32, ii, jj [E] i32,
    idx [E,N] i32, valid [E,N] bool, Q [E,N] f32, K4 (fx,fy,cx,cy) f32, img_size (W,H))."""
    rng = np.random.default_rng(5000 + seed)
    vv, uu = np.meshgrid(np.arange(h, dtype=np.float64), np.arange(w, dtype=np.float64),
                         indexing="ij")
    P = _surface(uu, vv, h, w).reshape(-1, 3)
    if num_pts is not None and num_pts < P.shape[0]:
        sel = np.linspace(0, P.shape[0] - 1, num_pts).astype(np.int64)
        P = P[sel]
    n = P.shape[0]
    Twc_gt = np.zeros((num_kf, 8))
    Twc = np.zeros((num_kf, 8))
    Xs = np.empty((num_kf, n, 3), np.float32)
    shifts = np.zeros(num_kf, np.int64)
    for k in range(num_kf):
        if k == 0:
            t, q, s = np.zeros(3), np.array([0, 0, 0, 1.0]), 1.0
        else:
            t = 0.05 * rng.standard_normal(3)
            q = _quat_from_rotvec(0.03 * rng.standard_normal(3))
            s = 1.0 + 0.02 * rng.uniform(-1, 1)
            shifts[k] = int(rng.integers(-3 * w, 3 * w))
        Twc_gt[k] = np.concatenate([t, q, [s]])
        R = _rotmat(q)
        Pk = P[(np.arange(n) + shifts[k]) % n]
        Xk = (Pk - t) @ R / s  # R^T (P - t) / s
        Xs[k] = (Xk + noise * rng.standard_normal((n, 3))).astype(np.float32)
        if k == 0:
            Twc[k] = Twc_gt[k]
        else:
            dq = _quat_from_rotvec(sigma_w * rng.standard_normal(3))
            Twc[k] = np.concatenate([t + sigma_t * rng.standard_normal(3), _quat_mul(dq, q),
                                     [s * np.exp(sigma_s * rng.standard_normal())]])
    ii, jj = graph_edges(num_kf, num_loops, rng)
    e = len(ii)
    base = np.arange(n, dtype=np.int64)
    idx = np.empty((e, n), np.int32)
    for k in range(e):
        idx[k] = (base + shifts[jj[k]] - shifts[ii[k]]) % n
    valid = rng.random((e, n)) < valid_frac
    Q = rng.uniform(1.2, 4.0, size=(e, n)).astype(np.float32)
    Cs = rng.uniform(1.0, 11.0, size=(num_kf, n)).astype(np.float32)
    K4 = np.array(intrinsics(h, w), np.float32)
    return dict(Twc=Twc.astype(np.float32), Twc_gt=Twc_gt.astype(np.float32), Xs=Xs, Cs=Cs, ii=ii,
                jj=jj, idx=idx, valid=valid, Q=Q, K4=K4, img_size=(w, h))


def make_track_pair(seed: int = 0, h: int = 384, w: int = 512, num_pts: int | None = None,
                    noise: float = 1e-3, valid_frac: float = 0.85):
    """Tracking problem (tracker.py:51-180): keyframe = KF 0 of `make_graph` (exact pose, its points
    sit on its own pixel grid), frame = KF 1 (perturbed pose, points in shifted order).  Returns
    dict(Xf, Xk [N,3], idx_f2k [N] i32 (keyframe pixel n <-> frame point idx_f2k[n]), valid_match
    [N,1] bool, Cf, Ck [N], Qff, Qkf [N], T_WCf (perturbed), T_WCf_gt, T_WCk [8], K4, img_size (h,w))."""
    g = make_graph(seed, 2, 0, h, w, num_pts, noise, valid_frac=valid_frac)
    assert (g["ii"][1], g["jj"][1]) == (1, 0)
    rng = np.random.default_rng(9000 + seed)
    n = g["Xs"].shape[1]
    return dict(Xf=g["Xs"][1], Xk=g["Xs"][0], idx_f2k=g["idx"][1], valid_match=g["valid"][1][:, None],
                Cf=g["Cs"][1], Ck=g["Cs"][0], Qff=rng.uniform(1.2, 4.0, n).astype(np.float32),
                Qkf=rng.uniform(1.2, 4.0, n).astype(np.float32), T_WCf=g["Twc"][1],
                T_WCf_gt=g["Twc_gt"][1], T_WCk=g["Twc"][0], K4=g["K4"], img_size=(h, w))


# ---------------------------------------------------------------------------------------------
# value distributions of the reference's own benchmark generators (incoherent worst cases)
# ---------------------------------------------------------------------------------------------
def reference_like_iter_proj(seed, b, h, w, n):
    """Distributions of benchmark_all_kernels.py:67-76: N(0,1) ray image, unit targets, uniform p."""
    rng = np.random.default_rng(seed)
    rays = rng.standard_normal((b, h, w, 9)).astype(np.float32)
    pts = rng.standard_normal((b, n, 3))
    pts = (pts / np.linalg.norm(pts, axis=-1, keepdims=True)).astype(np.float32)
    p0 = np.stack([rng.random((b, n)) * (w - 1), rng.random((b, n)) * (h - 1)], -1).astype(np.float32)
    return rays, pts, p0


def reference_like_refine(seed, b, h, w, d, n):
    """Distributions of benchmark_all_kernels.py:385-395: unit-norm random descriptors."""
    rng = np.random.default_rng(seed)
    D11 = rng.standard_normal((b, h, w, d))
    D11 = (D11 / np.linalg.norm(D11, axis=-1, keepdims=True)).astype(np.float32)
    D21 = rng.standard_normal((b, n, d))
    D21 = (D21 / np.linalg.norm(D21, axis=-1, keepdims=True)).astype(np.float32)
    p1 = np.stack([rng.integers(0, w, (b, n)), rng.integers(0, h, (b, n))], -1).astype(np.int32)
    return D11, D21, p1


def reference_like_gn(seed, num_kf=10, num_pts=500, num_edges=15):
    """Distributions of benchmark_all_kernels.py:16-42: random poses, permutation matches."""
    rng = np.random.default_rng(seed)
    Twc = np.zeros((num_kf, 8), np.float32)
    for i in range(num_kf):
        Twc[i, :3] = rng.standard_normal(3) * 0.5
        q = rng.standard_normal(4)
        Twc[i, 3:7] = q / np.linalg.norm(q)
        Twc[i, 7] = 1.0 + rng.random() * 0.1
    Xs = rng.standard_normal((num_kf, num_pts, 3)).astype(np.float32)
    Cs = (rng.random((num_kf, num_pts)) * 10 + 1).astype(np.float32)
    ii = rng.integers(0, num_kf, num_edges).astype(np.int32)
    jj = rng.integers(0, num_kf, num_edges).astype(np.int32)
    for e in range(num_edges):
        while jj[e] == ii[e]:
            jj[e] = rng.integers(0, num_kf)
    idx = np.stack([rng.permutation(num_pts) for _ in range(num_edges)]).astype(np.int32)
    valid = rng.random((num_edges, num_pts)) > 0.3
    Q = (rng.random((num_edges, num_pts)) * 3 + 1).astype(np.float32)
    return Twc, Xs, Cs, ii, jj, idx, valid, Q


# ---------------------------------------------------------------------------------------------
# device-side generators for full-size benchmark workloads (same scene model, torch on the GPU)
# ---------------------------------------------------------------------------------------------
def make_pairs_torch(seeds, device, h=384, w=512, d=24, desc="dyadic", noise=1e-3):
    """Config 2 (64 pairs): dict of CUDA tensors X11, X21 [B,H,W,3], D11, D21 [B,H,W,d], float32.
    Same scene model as make_pair; generated on the device so full-size batches take seconds."""
    import torch
    fx, fy, cx0, cy0 = intrinsics(h, w)
    vv, uu = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=device),
                            torch.arange(w, dtype=torch.float64, device=device), indexing="ij")

    def surface(u, v):
        z = 2.0 + 0.5 * torch.sin(2 * np.pi * u / w) * torch.cos(2 * np.pi * v / h) + 0.3 * (u / w)
        return torch.stack([(u - cx0) / fx * z, (v - cy0) / fy * z, z], -1)

    out = {k: [] for k in ("X11", "X21", "D11", "D21")}
    cell = 6
    gh, gw = h // cell + 3, w // cell + 3
    for seed in seeds:
        gen = torch.Generator(device=device)
        gen.manual_seed(1000 + int(seed))
        rng = np.random.default_rng(1000 + int(seed))
        s = 1.0 + 0.01 * rng.uniform(-1, 1)
        th = 0.01 * rng.uniform(-1, 1)
        t = rng.uniform(-3.0, 3.0, size=2)
        cx, cy = (w - 1) / 2.0, (h - 1) / 2.0
        du, dv = uu - cx, vv - cy
        u2 = cx + s * (np.cos(th) * du - np.sin(th) * dv) + t[0]
        v2 = cy + s * (np.sin(th) * du + np.cos(th) * dv) + t[1]
        X11 = surface(uu, vv)
        X21 = surface(u2, v2) + noise * torch.randn((h, w, 3), generator=gen, device=device, dtype=torch.float64)
        coarse = torch.randn((gh, gw, d), generator=gen, device=device, dtype=torch.float64)

        def field(u, v):
            gu = torch.clamp(u / cell + 1.0, 0, gw - 1.001)
            gv = torch.clamp(v / cell + 1.0, 0, gh - 1.001)
            x0, y0 = gu.floor().long(), gv.floor().long()
            fx_, fy_ = (gu - x0)[..., None], (gv - y0)[..., None]
            return ((1 - fx_) * (1 - fy_) * coarse[y0, x0] + (1 - fx_) * fy_ * coarse[y0 + 1, x0]
                    + fx_ * (1 - fy_) * coarse[y0, x0 + 1] + fx_ * fy_ * coarse[y0 + 1, x0 + 1])

        D11 = field(uu, vv) + 0.15 * torch.randn((h, w, d), generator=gen, device=device, dtype=torch.float64)
        D21 = field(u2, v2) + 0.15 * torch.randn((h, w, d), generator=gen, device=device, dtype=torch.float64)
        if desc == "dyadic":
            D11 = torch.clamp(torch.round(D11 * 64.0) / 64.0, -2.0, 2.0)
            D21 = torch.clamp(torch.round(D21 * 64.0) / 64.0, -2.0, 2.0)
        else:
            D11 = D11 / D11.norm(dim=-1, keepdim=True)
            D21 = D21 / D21.norm(dim=-1, keepdim=True)
        for k, v in (("X11", X11), ("X21", X21), ("D11", D11), ("D21", D21)):
            out[k].append(v.float())
    return {k: torch.stack(v).contiguous() for k, v in out.items()}


def make_graph_torch(device, seed=0, num_kf=200, num_loops=103, h=384, w=512, noise=1e-3,
                     sigma_t=0.02, sigma_w=0.01, sigma_s=0.01, valid_frac=0.85, edge_range=None):
    """Config 3 / 5 at full num_pts, generated on the device.  Same model as make_graph (poses and
    topology come from the same numpy stream, so every rank builds the identical graph).
    `edge_range=(b, e)`: only materialise idx/valid/Q rows of directed edges [b, e) (edge-sharded
    ranks), other rows are left out; ii/jj always cover the whole graph.
    Returns dict(Twc, Twc_gt [K,8] f32 numpy; ii, jj numpy; Xs [K,N,3], Cs [K,N], idx, valid (uint8),
    Q CUDA tensors; K4; img_size)."""
    import torch
    rng = np.random.default_rng(5000 + seed)
    fx, fy, cx0, cy0 = intrinsics(h, w)
    vv, uu = torch.meshgrid(torch.arange(h, dtype=torch.float64, device=device),
                            torch.arange(w, dtype=torch.float64, device=device), indexing="ij")
    z = 2.0 + 0.5 * torch.sin(2 * np.pi * uu / w) * torch.cos(2 * np.pi * vv / h) + 0.3 * (uu / w)
    P = torch.stack([(uu - cx0) / fx * z, (vv - cy0) / fy * z, z], -1).reshape(-1, 3)
    n = P.shape[0]
    gen = torch.Generator(device=device)
    gen.manual_seed(5000 + seed)
    Twc_gt = np.zeros((num_kf, 8))
    Twc = np.zeros((num_kf, 8))
    Xs = torch.empty((num_kf, n, 3), dtype=torch.float32, device=device)
    shifts = np.zeros(num_kf, np.int64)
    base = torch.arange(n, device=device)
    for k in range(num_kf):
        if k == 0:
            t, q, s = np.zeros(3), np.array([0, 0, 0, 1.0]), 1.0
        else:
            t = 0.05 * rng.standard_normal(3)
            q = _quat_from_rotvec(0.03 * rng.standard_normal(3))
            s = 1.0 + 0.02 * rng.uniform(-1, 1)
            shifts[k] = int(rng.integers(-3 * w, 3 * w))
        Twc_gt[k] = np.concatenate([t, q, [s]])
        R = torch.from_numpy(_rotmat(q)).to(device)
        Pk = P[(base + int(shifts[k])) % n]
        Xk = (Pk - torch.from_numpy(t).to(device)) @ R / s
        Xs[k] = (Xk + noise * torch.randn((n, 3), generator=gen, device=device, dtype=torch.float64)).float()
        if k == 0:
            Twc[k] = Twc_gt[k]
        else:
            dq = _quat_from_rotvec(sigma_w * rng.standard_normal(3))
            Twc[k] = np.concatenate([t + sigma_t * rng.standard_normal(3), _quat_mul(dq, q),
                                     [s * np.exp(sigma_s * rng.standard_normal())]])
    ii, jj = graph_edges(num_kf, num_loops, rng)
    e = len(ii)
    eb, ee = (0, e) if edge_range is None else edge_range
    ne = ee - eb
    idx = torch.empty((ne, n), dtype=torch.int32, device=device)
    for r, k in enumerate(range(eb, ee)):
        idx[r] = ((base + int(shifts[jj[k]] - shifts[ii[k]])) % n).int()
    Cs = 1.0 + 10.0 * torch.rand((num_kf, n), generator=gen, device=device)
    # per-edge seeds: a rank that materialises only its edge rows sees the same rows as a full build
    valid = torch.empty((ne, n), dtype=torch.uint8, device=device)
    Q = torch.empty((ne, n), dtype=torch.float32, device=device)
    for r, k in enumerate(range(eb, ee)):
        gen.manual_seed(900000 + 1000 * seed + k)
        valid[r] = (torch.rand(n, generator=gen, device=device) < valid_frac).to(torch.uint8)
        Q[r] = 1.2 + 2.8 * torch.rand(n, generator=gen, device=device)
    return dict(Twc=Twc.astype(np.float32), Twc_gt=Twc_gt.astype(np.float32), Xs=Xs, Cs=Cs, ii=ii, jj=jj,
                idx=idx, valid=valid, Q=Q, K4=np.array([fx, fy, cx0, cy0], np.float32), img_size=(w, h),
                edge_range=(eb, ee))
