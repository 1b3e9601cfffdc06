"""GPU parity tests for the matching stages (run on the B200 box: pytest -m gpu).

Every call goes through the C ABI (ctypes -> libm3s_b200.so).  Results are compared with the golden
vectors produced by the reference itself and with the CPU oracle on seeded inputs.  The bar is
bit-exact: positions (float32), validity and integer indices must be identical.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def K():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from mast3r_slam_b200 import kernels
    assert kernels.is_available(), "CUDA extension not loaded"
    return kernels


@pytest.fixture(scope="module")
def synth():
    from mast3r_slam_b200 import synth as s
    return s


def test_iter_proj_golden_random(K, golden):
    g = golden("iter_proj_random")
    p, v = K.iter_proj(g["rays"], g["pts"], g["p_init"], 10, 1e-8, 1e-6)
    assert p.dtype == np.float32 and v.dtype == np.bool_
    assert np.array_equal(v, g["valid"])
    assert np.array_equal(p, g["p"])


@pytest.mark.parametrize("it", [1, 3, 10])
def test_iter_proj_golden_scene(K, golden, it):
    g = golden("iter_proj_scene")
    p, v = K.iter_proj(g["rays"], g["pts"], g["p_init"], it, 1e-8, 1e-6)
    assert np.array_equal(v, g[f"valid_{it}"])
    assert np.array_equal(p, g[f"p_{it}"])


def test_iter_proj_global_break(K, golden):
    """kernels.py:239-241: the loop stops for ALL points once max ||delta|| < thresh."""
    g = golden("iter_proj_break")
    p, v = K.iter_proj(g["rays_same"], g["pts_same"], g["p0_same"], 10, 1e-8, 1e-6)
    assert np.array_equal(p, g["p_same"]) and np.array_equal(v, g["valid_same"])
    p, v = K.iter_proj(g["rays"], g["pts"], g["p_init"], 10, 1e-8, float(g["thresh"]))
    assert np.array_equal(p, g["p"]) and np.array_equal(v, g["valid"])
    # same inputs, tiny threshold: must NOT stop early (differs from the early-stopped result)
    p10, _ = K.iter_proj(g["rays"], g["pts"], g["p_init"], 10, 1e-8, 0.0)
    assert not np.array_equal(p10, g["p"])


@pytest.mark.parametrize("shape", [(1, 33, 47, 1000), (3, 24, 40, 777), (2, 64, 64, 4096)])
def test_iter_proj_vs_oracle_random(K, oracle, synth, shape):
    b, h, w, n = shape
    rays, pts, p0 = synth.reference_like_iter_proj(100 + n, b, h, w, n)
    p0[:, :5, 0] -= 7.0
    p0[:, 5:9, 1] += h
    for it, lam in ((10, 1e-8), (4, 1e-3)):
        p, v = K.iter_proj(rays, pts, p0, it, lam, 1e-6)
        po, vo = oracle.iter_proj(rays, pts, p0, it, lam, 1e-6)
        assert np.array_equal(v, vo)
        assert np.array_equal(p, po)


def test_iter_proj_edge_cases(K, oracle, synth):
    rays, pts, p0 = synth.reference_like_iter_proj(5, 1, 16, 16, 64)
    # zero iterations: clamp + validity of the initial positions only
    p, v = K.iter_proj(rays, pts, p0, 0, 1e-8, 1e-6)
    po, vo = oracle.iter_proj(rays, pts, p0, 0, 1e-8, 1e-6)
    assert np.array_equal(p, po) and np.array_equal(v, vo)
    # empty point set
    p, v = K.iter_proj(rays, pts[:, :0], p0[:, :0], 10)
    assert p.shape == (1, 0, 2) and v.shape == (1, 0)
    # degenerate gradients (constant image): det clamps to 1e-10
    flat = np.ones_like(rays)
    p, v = K.iter_proj(flat, pts, p0, 3, 1e-8, 1e-6)
    po, vo = oracle.iter_proj(flat, pts, p0, 3, 1e-8, 1e-6)
    assert np.array_equal(p, po) and np.array_equal(v, vo)
    # torch CUDA tensors in -> torch CUDA tensors out, same bits
    import torch
    pt, vt = K.iter_proj(torch.from_numpy(rays).cuda(), torch.from_numpy(pts).cuda(),
                         torch.from_numpy(p0).cuda(), 10)
    assert pt.is_cuda and vt.dtype == torch.bool
    po, vo = oracle.iter_proj(rays, pts, p0, 10)
    assert np.array_equal(pt.cpu().numpy(), po) and np.array_equal(vt.cpu().numpy(), vo)


@pytest.mark.parametrize("impl", ["quad", "tile", "cache", "generic"])
@pytest.mark.parametrize("hw", [(48, 64), (37, 52), (16, 36), (50, 70)])
def test_iter_proj_full_frame_impls(K, oracle, synth, monkeypatch, impl, hw):
    """Full-frame queries (N == H*W) through each implementation (2x2-points-per-thread quad kernel
    with its eviction queue, one-point tile kernel, generic kernel): coherent flow, incoherent
    starts (evictions / global fallback), starts outside the image, NaN starts, ragged sizes
    (partial tiles, odd H, W % 4 != 0 falls back by itself), several iteration counts, B = 2."""
    monkeypatch.setenv("M3S_ITER_PROJ_IMPL", impl)
    h, w = hw
    b = 2
    pairs = [synth.make_pair(7 + i, h, w, 24, "dyadic") for i in range(b)]
    X11 = np.stack([q["X11"] for q in pairs])
    X21 = np.stack([q["X21"] for q in pairs])
    rays, pts, p0 = oracle.prep_for_iter_proj(X11, X21, None)
    rng = np.random.default_rng(h * 1000 + w)
    starts = {"identity": p0}
    jit = p0 + rng.uniform(-1.5, 1.5, p0.shape).astype(np.float32)
    starts["jitter"] = jit
    wild = jit.copy()
    sel = rng.random(p0.shape[:2]) < 0.05
    wild[sel] = np.stack([rng.uniform(-6, w + 6, sel.sum()), rng.uniform(-6, h + 6, sel.sum())], -1)
    wild[0, 3] = np.nan
    wild[1, 10, 1] = np.nan
    starts["wild"] = wild.astype(np.float32)
    starts["random"] = np.stack([rng.uniform(0, w, p0.shape[:2]), rng.uniform(0, h, p0.shape[:2])], -1).astype(np.float32)
    for name, st in starts.items():
        for it, lam in ((10, 1e-8), (3, 1e-4), (1, 1e-8)):
            p, v = K.iter_proj(rays, pts, st, it, lam, 1e-6)
            if name == "wild":  # numpy cannot index with NaN positions: compare the finite points
                fin = np.isfinite(st).all(-1)
                st_o = np.where(fin[..., None], st, 0.0).astype(np.float32)
                po, vo = oracle.iter_proj(rays, pts, st_o, it, lam, 1e-6)
                assert np.array_equal(p[fin], po[fin]) and np.array_equal(v[fin], vo[fin]), (name, it)
                assert not v[~fin].any()
            else:
                po, vo = oracle.iter_proj(rays, pts, st, it, lam, 1e-6)
                assert np.array_equal(v, vo), (name, it)
                assert np.array_equal(p, po), (name, it, int((p != po).any(-1).sum()))
    # the reference's own random ray images (incoherent gradients: large steps, many evictions)
    rr, pp, q0 = synth.reference_like_iter_proj(3, b, h, w, h * w)
    p, v = K.iter_proj(rr, pp, q0, 10, 1e-8, 1e-6)
    po, vo = oracle.iter_proj(rr, pp, q0, 10, 1e-8, 1e-6)
    assert np.array_equal(p, po) and np.array_equal(v, vo)


def test_iter_proj_global_break_full_frame(K, oracle, synth, monkeypatch):
    """global early exit (kernels.py:239-241) through the quad kernel and its queue: a threshold
    between two iterations' maxima must stop every point at the same iteration as the oracle."""
    h, w = 32, 64
    pair = synth.make_pair(11, h, w, 24, "dyadic")
    rays, pts, p0 = oracle.prep_for_iter_proj(pair["X11"][None], pair["X21"][None], None)
    p0 = p0 + np.random.default_rng(0).uniform(-0.4, 0.4, p0.shape).astype(np.float32)
    ref = [oracle.iter_proj(rays, pts, p0, k, 1e-8, 0.0)[0] for k in range(0, 9)]
    step = [float(np.sqrt(((ref[k + 1] - ref[k]).astype(np.float32) ** 2).sum(-1)).max()) for k in range(8)]
    for impl in ("quad", "tile", "cache", "generic"):
        monkeypatch.setenv("M3S_ITER_PROJ_IMPL", impl)
        for k in (2, 5):
            thr = 0.5 * (step[k] + step[k - 1]) if step[k] < step[k - 1] else step[k] * 1.01
            p, v = K.iter_proj(rays, pts, p0, 8, 1e-8, thr)
            po, vo = oracle.iter_proj(rays, pts, p0, 8, 1e-8, thr)
            assert np.array_equal(p, po) and np.array_equal(v, vo), (impl, k)


def test_prep_and_tail_vs_oracle(K, oracle, synth):
    pair = synth.make_pair(2, 40, 56)
    X11, X21 = pair["X11"][None], pair["X21"][None]
    X11b = np.concatenate([X11, X11[:, ::-1].copy()])
    X21b = np.concatenate([X21, X21[:, :, ::-1].copy()])
    idx0 = np.random.default_rng(0).integers(0, 40 * 56, (2, 40 * 56)).astype(np.int32)
    for idx in (None, idx0):
        got = K.prep_for_iter_proj(X11b, X21b, idx)
        ref = oracle.prep_for_iter_proj(X11b, X21b, idx)
        for a, b in zip(got, ref):
            assert np.array_equal(a, b)
    rng = np.random.default_rng(1)
    p1 = np.stack([rng.integers(-3, 60, (2, 40 * 56)), rng.integers(-3, 44, (2, 40 * 56))], -1).astype(np.int32)
    vp = rng.random((2, 40 * 56)) < 0.8
    idx, valid = K.match_tail(X11b, X21b, p1, vp, 0.1)
    idx_o, valid_o = oracle.match_tail(X11b, X21b, p1, vp, 0.1)
    assert idx.dtype == np.int32 and valid.shape == (2, 40 * 56, 1)
    assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)


@pytest.mark.parametrize("hw", [(40, 56), (8, 32), (9, 36), (37, 52), (3, 4), (1, 8), (50, 70), (16, 33)])
def test_prep_shapes_vs_oracle(K, oracle, synth, hw):
    """prep through the tiled kernel (32 x 8 tiles, shared halo of unit rays, float4 row stores):
    full and partial tiles in x and y, single-row / 4-pixel-wide frames, and W % 4 != 0 shapes that
    take the one-thread-per-pixel kernel -- all bit-identical to the oracle, B = 2, with and without
    idx_init (incl. negative indices: python floor semantics)."""
    h, w = hw
    rng = np.random.default_rng(h * 100 + w)
    X11 = rng.normal(size=(2, h, w, 3)).astype(np.float32)
    X11[..., 2] += 3.0
    X21 = rng.normal(size=(2, h * w, 3)).astype(np.float32)
    X21[0, 0] = 0.0   # zero vector: 0 / sqrt(1e-10)
    idx0 = rng.integers(-h * w, h * w, (2, h * w)).astype(np.int32)
    for idx in (None, idx0):
        got = K.prep_for_iter_proj(X11, X21.reshape(2, h, w, 3), idx)
        ref = oracle.prep_for_iter_proj(X11, X21.reshape(2, h, w, 3), idx)
        for a, b in zip(got, ref):
            assert a.shape == b.shape and np.array_equal(a, b)


@pytest.mark.parametrize("key", ["r3_d0", "r3_d2", "r1_d1", "r2_d3"])
def test_refine_golden_dyadic(K, golden, key):
    g = golden("refine_dyadic")
    r, d = int(key[1]), int(key[4])
    out = K.refine_matches(g["D11"], g["D21"], g["p1"], r, d)
    assert out.dtype == np.int32
    assert np.array_equal(out, g[key])


def _near_tie_ok(D11, D21, a, b):
    """rows where the two answers differ must be rounding-level ties of the fp32 score"""
    bad = np.argwhere((a != b).any(-1))
    for bi, n in bad:
        q = D21[bi, n].astype(np.float64)
        sa = q @ D11[bi, a[bi, n, 1], a[bi, n, 0]].astype(np.float64)
        sb = q @ D11[bi, b[bi, n, 1], b[bi, n, 0]].astype(np.float64)
        assert abs(sa - sb) <= 4 * np.finfo(np.float32).eps * max(abs(sa), abs(sb), 1.0), (bi, n, sa, sb)
    return len(bad)


def test_refine_golden_unit(K, golden):
    g = golden("refine_unit")
    out = K.refine_matches(g["D11"], g["D21"], g["p1"], 3, 2)
    assert _near_tie_ok(g["D11"], g["D21"], out, g["r3_d2"]) <= 2
    gf = golden("refine_floatp")
    out = K.refine_matches(g["D11"], g["D21"][:, :50], gf["p1f"], 2, 0)
    assert np.array_equal(out, gf["r2_d0"])


@pytest.mark.parametrize("d", [24, 16, 32, 64, 20, 7])
def test_refine_vs_oracle_dims(K, oracle, d):
    """all descriptor widths (vectorised and generic kernels), dyadic values -> exact"""
    rng = np.random.default_rng(d)
    b, h, w, n = 2, 30, 45, 1200
    D11 = (rng.integers(-128, 129, (b, h, w, d)) / 64.0).astype(np.float32)
    D21 = (rng.integers(-128, 129, (b, n, d)) / 64.0).astype(np.float32)
    p1 = np.stack([rng.integers(-4, w + 4, (b, n)), rng.integers(-4, h + 4, (b, n))], -1).astype(np.int32)
    for r in (0, 1, 3, 5):
        out = K.refine_matches(D11, D21, p1, r, 0)
        ref = oracle.refine_matches(D11, D21, p1, r, 0)
        assert np.array_equal(out, ref), (d, r)
    # ties: constant descriptors -> the first candidate in (dy, dx) scan order wins
    D11c = np.ones_like(D11)
    out = K.refine_matches(D11c, D21, p1, 2, 0)
    ref = oracle.refine_matches(D11c, D21, p1, 2, 0)
    assert np.array_equal(out, ref)


@pytest.mark.parametrize("desc", ["dyadic", "unit"])
def test_match_pipeline_vs_oracle(K, oracle, synth, desc):
    """matching.py:339-461 end to end in one C-ABI call, B=2"""
    ps = synth.make_pairs([7, 8], 72, 96, 24, desc)
    idx, valid = K.match_iterative_proj_device(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    idx_o, valid_o = oracle.match_iterative_proj(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    assert idx.shape == (2, 72 * 96) and valid.shape == (2, 72 * 96, 1)
    if desc == "dyadic":
        assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)
    else:
        # non-dyadic scores: summation order (FMA chain vs the oracle's separate ops) may flip
        # rounding-level ties only
        same = idx == idx_o
        assert (~same).mean() < 1e-3
        assert np.array_equal(valid[same], valid_o[same])
    # with an initial correspondence and without refinement
    idx0 = idx_o.copy()
    a = K.match_iterative_proj_device(ps["X11"], ps["X21"], ps["D11"], ps["D21"], idx0, radius=0)
    b = oracle.match_iterative_proj(ps["X11"], ps["X21"], ps["D11"], ps["D21"], idx0, radius=0)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_match_host_pipeline(K, oracle, synth):
    """m3s_match_iterative_proj_host: chunked upload / compute / download on rotating streams gives
    the same result as the oracle for ragged chunking, with an initial correspondence, without
    refinement, with pinned torch inputs and caller-provided pinned outputs"""
    import torch
    ps = synth.make_pairs([11, 12, 13, 14, 15, 16, 17], 48, 64, 24, "dyadic")
    a = (ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    idx_o, valid_o = oracle.match_iterative_proj(*a)
    for chunk in (1, 2, 3, 7, 64):
        idx, valid = K.match_iterative_proj_host(*a, chunk_pairs=chunk)
        assert isinstance(idx, np.ndarray) and idx.dtype == np.int32 and valid.dtype == np.bool_
        assert valid.shape == (7, 48 * 64, 1)
        assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o), chunk
    # numpy inputs to the generic entry point take the same route
    idx, valid = K.match_iterative_proj_device(*a)
    assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)
    # initial correspondence, no refinement
    r = oracle.match_iterative_proj(*a, idx_o.copy(), radius=0)
    g = K.match_iterative_proj_host(*a, idx_o.copy(), radius=0, chunk_pairs=2)
    assert np.array_equal(g[0], r[0]) and np.array_equal(g[1], r[1])
    # pinned torch tensors in, caller-provided pinned outputs
    pin = [torch.from_numpy(x).pin_memory() for x in a]
    out_idx = torch.empty((7, 48 * 64), dtype=torch.int32).pin_memory()
    out_valid = torch.empty((7, 48 * 64), dtype=torch.uint8).pin_memory()
    ti, tv = K.match_iterative_proj_host(*pin, chunk_pairs=2, out=(out_idx, out_valid))
    assert isinstance(ti, torch.Tensor) and ti.data_ptr() == out_idx.data_ptr()
    assert np.array_equal(ti.numpy(), idx_o) and np.array_equal(tv.numpy(), valid_o)
    # empty batch
    e = K.match_iterative_proj_host(*(x[:0] for x in a))
    assert e[0].shape == (0, 48 * 64)


def test_l2_match_entry_point(K, oracle, synth):
    from mast3r_slam_b200 import matching
    from mast3r_slam_b200.config import set_config
    ps = synth.make_pairs([3], 48, 64)
    # default configuration (config.py:65 use_simple=True): the match_simple route
    idx, valid = matching.match(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    idx_s, valid_s = oracle.match_simple(ps["X11"], ps["X21"], None)
    assert np.array_equal(idx, idx_s) and np.array_equal(valid, valid_s) and valid.shape == valid_s.shape
    set_config({"matching": {"use_simple": False}})
    try:
        idx, valid = matching.match(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    finally:
        set_config({})
    idx_o, valid_o = oracle.match_iterative_proj(ps["X11"], ps["X21"], ps["D11"], ps["D21"],
                                                 radius=3, dilation_max=2)
    assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)


def test_full_size_config1(K, oracle, synth):
    """BASELINE.json config 1: one 512x384 pair, D=24, 10 iterations, radius 3 -- bit-exact against
    the oracle at full size, plus size-independent properties."""
    pair = synth.make_pair(0, 384, 512, 24, "dyadic")
    X11, X21, D11, D21 = (pair[k][None] for k in ("X11", "X21", "D11", "D21"))
    rays, pts, p0 = K.prep_for_iter_proj(X11, X21, None)
    ro, po_, p0o = oracle.prep_for_iter_proj(X11, X21, None)
    assert np.array_equal(rays, ro) and np.array_equal(pts, po_) and np.array_equal(p0, p0o)
    p, v = K.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    po, vo = oracle.iter_proj(rays, pts, p0, 10, 1e-8, 1e-6)
    assert np.array_equal(p, po) and np.array_equal(v, vo)
    p1 = K.refine_matches(D11.reshape(1, 384, 512, 24), D21.reshape(1, -1, 24), p.astype(np.int32), 3, 0)
    p1o = oracle.refine_matches(D11, D21.reshape(1, -1, 24), p.astype(np.int32), 3, 0)
    assert np.array_equal(p1, p1o)
    idx, valid = K.match_iterative_proj_device(X11, X21, D11, D21)
    idx_o, valid_o = oracle.match_iterative_proj(X11, X21, D11, D21)
    assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)
    # properties: refined positions stay inside the window and inside the image; idempotent centre
    c = p.astype(np.int32)
    assert (np.abs(p1 - c) <= 3).all()
    assert (p1[..., 0] >= 0).all() and (p1[..., 0] < 512).all() and (p1[..., 1] < 384).all()
    # most matches recover the true flow to within the refinement window
    flow = pair["flow_uv"].reshape(-1, 2)
    ok = valid[0, :, 0]
    err = np.abs(p[0][ok] - flow[ok]).max(-1)
    assert np.median(err) < 0.5 and ok.mean() > 0.9


@pytest.mark.parametrize("hw", [(50, 70), (16, 32), (37, 129), (16, 36)])
def test_refine_tiled_path(K, oracle, synth, hw):
    """full-frame queries (N == H*W, D == 24): the shared-memory tiled kernel, coherent and
    incoherent centres, ragged image sizes, all radii it accepts, B = 2"""
    h, w = hw
    rng = np.random.default_rng(h * w)
    b, d = 2, 24
    D11 = (rng.integers(-128, 129, (b, h, w, d)) / 64.0).astype(np.float32)
    D21 = (rng.integers(-128, 129, (b, h * w, d)) / 64.0).astype(np.float32)
    yy, xx = np.meshgrid(np.arange(h), np.arange(w), indexing="ij")
    ident = np.stack([xx, yy], -1).reshape(1, h * w, 2)
    coherent = np.repeat(ident, b, 0) + np.array([2, -1])
    jitter = coherent + rng.integers(-1, 2, coherent.shape)
    # a few wild outliers inside otherwise coherent tiles + centres outside the image
    outl = jitter.copy()
    sel = rng.random((b, h * w)) < 0.03
    outl[sel] = np.stack([rng.integers(-5, w + 5, sel.sum()), rng.integers(-5, h + 5, sel.sum())], -1)
    random = np.stack([rng.integers(-4, w + 4, (b, h * w)), rng.integers(-4, h + 4, (b, h * w))], -1)
    # stacked queries of one thread > 25 px apart in x but all inside the 40-px staged window: their
    # column masks do not fit 32 bits, the far ones must take the global path (same result)
    spread = np.repeat(ident, b, 0).copy()
    spread[..., 0] = spread[..., 0] % 4 + np.where(spread[..., 1] % 4 == 2, 28, 0)
    for name, p1 in (("coherent", coherent), ("jitter", jitter), ("outliers", outl), ("random", random),
                     ("spread", spread)):
        p1 = p1.astype(np.int32)
        for r in (1, 2, 3):
            out = K.refine_matches(D11, D21, p1, r, 0)
            ref = oracle.refine_matches(D11, D21, p1, r, 0)
            assert np.array_equal(out, ref), (name, r, int((out != ref).any(-1).sum()))
    # float positions (truncation fused into the load) through the tiled kernel
    pf = (jitter + rng.random(jitter.shape) * 0.99).astype(np.float32)
    out = K.refine_matches(D11, D21, pf, 3, 0)
    ref = oracle.refine_matches(D11, D21, pf.astype(np.int32), 3, 0)
    assert np.array_equal(out, ref)
    # ties everywhere: first candidate in scan order must win in the tiled kernel too
    out = K.refine_matches(np.ones_like(D11), D21, jitter.astype(np.int32), 3, 0)
    ref = oracle.refine_matches(np.ones_like(D11), D21, jitter.astype(np.int32), 3, 0)
    assert np.array_equal(out, ref)


def test_full_size_config2_batch_properties(K, oracle, synth):
    """BASELINE.json config 2: a multi-pair batch at 512x384 in one call.  The whole batch is
    checked through size-independent properties; 2 of its pairs bit-exactly against the oracle;
    and batching itself must not change any pair's result (pairs are independent)."""
    import torch
    B = 8
    ps = synth.make_pairs_torch(range(B), torch.device("cuda", 0), 384, 512, 24, "dyadic")
    idx, valid = K.match_iterative_proj_device(ps["X11"], ps["X21"], ps["D11"], ps["D21"])
    idx, valid = idx.cpu().numpy(), valid.cpu().numpy()
    assert idx.shape == (B, 196608) and valid.shape == (B, 196608, 1)
    x, y = idx % 512, idx // 512
    ok = valid[..., 0]
    assert ((x[ok] >= 0) & (x[ok] < 512) & (y[ok] >= 0) & (y[ok] < 384)).all()
    assert ok.mean() > 0.9
    for b in (0, B - 1):
        one = {k: v[b:b + 1].cpu().numpy() for k, v in ps.items()}
        idx_o, valid_o = oracle.match_iterative_proj(one["X11"], one["X21"], one["D11"], one["D21"])
        assert np.array_equal(idx[b:b + 1], idx_o) and np.array_equal(valid[b:b + 1], valid_o)
        i1, v1 = K.match_iterative_proj_device(ps["X11"][b:b + 1], ps["X21"][b:b + 1], ps["D11"][b:b + 1],
                                               ps["D21"][b:b + 1])
        assert np.array_equal(i1.cpu().numpy(), idx[b:b + 1]) and np.array_equal(v1.cpu().numpy(), valid[b:b + 1])


@pytest.mark.parametrize("hw", [(48, 64), (37, 52), (50, 70)])
def test_match_fused_prep_path(K, oracle, synth, monkeypatch, hw):
    """M3S_MATCH_FUSED=1: prep_for_iter_proj computed inside the iter_proj tile kernel (unit rays of
    the window halo in shared memory, gradients as differences, no ray image in HBM): the whole
    match_iterative_proj must stay bit-identical to the oracle, with and without idx_init, also for
    partial tiles, W % 4 != 0 and starts that leave the window (global fallback from X11)."""
    monkeypatch.setenv("M3S_MATCH_FUSED", "1")
    h, w = hw
    ps = [synth.make_pair(31 + i, h, w, 24, "dyadic") for i in range(2)]
    a = [np.stack([q[k] for q in ps]) for k in ("X11", "X21", "D11", "D21")]
    idx, valid = K.match_iterative_proj_device(*a)
    idx_o, valid_o = oracle.match_iterative_proj(*a)
    assert np.array_equal(idx, idx_o) and np.array_equal(valid, valid_o)
    rng = np.random.default_rng(h)
    idx0 = rng.integers(-h * w, h * w, (2, h * w)).astype(np.int32)   # incoherent starts: fallback path
    g = K.match_iterative_proj_device(*a, idx0)
    r = oracle.match_iterative_proj(*a, idx0)
    assert np.array_equal(g[0], r[0]) and np.array_equal(g[1], r[1])


def test_iter_proj_very_wide_image(K, oracle):
    """W large enough that f32(W - 1.001) == W - 1: the upper clip bound no longer keeps x0 <= W - 2,
    so the four corners need the reference's min(x0 + 1, W - 1) (kernels.py:178-180).  Full-frame
    query layout (N == H*W), i.e. the shape that would otherwise select the tile kernel."""
    h, w = 2, 40000
    assert np.float32(w - 1.001) == np.float32(w - 1)
    rng = np.random.default_rng(5)
    rays = rng.normal(size=(1, h, w, 9)).astype(np.float32)
    rays[..., :3] /= np.linalg.norm(rays[..., :3], axis=-1, keepdims=True)
    rays[..., 3:] *= 0.05
    pts = rng.normal(size=(1, h * w, 3)).astype(np.float32)
    pts /= np.linalg.norm(pts, axis=-1, keepdims=True)
    p0 = np.stack([np.tile(np.arange(w, dtype=np.float32), h), np.repeat(np.arange(h, dtype=np.float32), w)], -1)[None]
    p0[0, -64:, 0] = w - 1  # starts in the last column: x0 = W - 1 after the clip
    p0[0, -32:, 0] = w + 3.5
    p, v = K.iter_proj(rays, pts, p0, 3, 1e-8, 1e-6)
    po, vo = oracle.iter_proj(rays, pts, p0, 3, 1e-8, 1e-6)
    assert np.array_equal(p, po) and np.array_equal(v, vo)


@pytest.mark.parametrize("impl", ["tile", "cache", "quad", "generic"])
def test_iter_proj_extreme_determinants(K, oracle, synth, monkeypatch, impl):
    """The LM step divides by det = a d - b^2 clamped at 1e-10 (kernels.py:217-222).  The kernels take the
    fast path of the IEEE division below 1e30 and the library division above; gradients scaled so that det
    lands on both sides of that guard, overflows to inf, or turns NaN must all give the reference's bits."""
    monkeypatch.setenv("M3S_ITER_PROJ_IMPL", impl)
    h, w = 32, 64
    pair = synth.make_pair(21, h, w, 24, "dyadic")
    rays, pts, p0 = oracle.prep_for_iter_proj(pair["X11"][None], pair["X21"][None], None)
    rng = np.random.default_rng(5)
    for scale in (1e3, 3e6, 1e8, 1e10, 3e10):   # det ~ scale^4 x 3e-11: 3e1 .. 2e31 (both sides of the guard)
        r = rays.copy()
        r[..., 3:] *= np.float32(scale)
        p, v = K.iter_proj(r, pts, p0, 4, 1e-8, 1e-6)
        with np.errstate(all="ignore"):
            po, vo = oracle.iter_proj(r, pts, p0, 4, 1e-8, 1e-6)
        fin = np.isfinite(po).all(-1)          # numpy cannot index with the NaN positions a blown-up step leaves
        assert np.array_equal(p[fin], po[fin]) and np.array_equal(v[fin], vo[fin]), scale
    # per-pixel mixture of scales in one image incl. det = inf and inf - inf = NaN; ONE iteration, so that no
    # NaN position reaches the sampler (numpy raises on those, the C oracle cannot index them either)
    r = rays.copy()
    mix = rng.choice(np.array([1.0, 1e4, 1e7, 3e9, 1e12, 1e19], np.float32), size=r.shape[:3])
    r[..., 3:] *= mix[..., None]
    p, v = K.iter_proj(r, pts, p0, 1, 1e-8, 1e-6)
    with np.errstate(all="ignore"):
        po, vo = oracle.iter_proj(r, pts, p0, 1, 1e-8, 1e-6)
    same = (p == po) | (np.isnan(p) & np.isnan(po))
    assert same.all() and np.array_equal(v, vo)
