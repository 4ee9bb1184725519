"""GPU parity, part 2: the C-ABI (through flagg.engine) against the oracle on seeded
inputs at sizes the oracle finishes in seconds -- every network size, the radix
fallback, ragged / unaligned segments, edge values -- plus size-independent properties at
BASELINE.json's full sizes."""
import numpy as np
import pytest
import torch

from conftest import assert_close_scaled, bits
from oracle import oracle64, ref_port

pytestmark = pytest.mark.gpu

EDGE = np.array([0.0, -0.0, 1.0, -1.0, 2.5, np.inf, -np.inf, 1e-38, -1e-38, 1e30], dtype=np.float32)


def rows_model(rng, n, d, byz=0):
    theta = rng.normal(0, 0.05, d)
    X = theta[None, :] + rng.normal(0, 1e-3, (n, d))
    for b in range(byz):
        X[b] = theta + 10 * (X[b] - theta)
    return X.astype(np.float32)


def dev_lists(X, seg=None):
    """Rows as separately allocated device tensors (list-of-lists, like run_experiments.py:238)."""
    n, d = X.shape
    seg = seg or [d]
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    return [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(seg))] for i in range(n)]


def run_select(X, mode, k=0, seg=None):
    from flagg import engine
    u = engine.Updates(dev_lists(X, seg))
    out = engine.colwise_select(u, mode, k).cpu().numpy()
    u.close()
    return out


@pytest.mark.parametrize("n", list(range(1, 41)) + [44, 47, 48, 50, 63, 64, 65, 71, 72, 96, 99, 100, 101, 127, 128])
def test_median_every_network_size(n):
    from flagg import MEDIAN
    rng = np.random.RandomState(1000 + n)
    d = 777
    X = rows_model(rng, n, d, byz=n // 5)
    X[:, :200] = rng.choice(EDGE, size=(n, 200))
    X[:, 200:232] = rng.choice(np.array([0.0, -0.0], dtype=np.float32), size=(n, 32))   # mixed-sign zero columns
    Xi = X.view(np.uint32)
    for j in range(232, 260):
        Xi[rng.randint(n), j] = rng.choice([0x7FC00000, 0xFFC00001, 0x7F800123])        # NaN payloads
    got = run_select(X, MEDIAN, seg=[300, 77, 400])
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X))), n


def dense_lists(X, seg, pitch_pad=4):
    """Rows as slices of ONE device matrix whose row pitch is a multiple of 16 bytes: the layout
    flagg_view_create recognises as dense and feeds through 2-D TMA tensor maps (tensors whose
    column offset is a multiple of 4 floats; otherwise the view is not 16-byte aligned and the
    LDG kernel runs)."""
    n, d = X.shape
    ld = (d + 3) // 4 * 4 + pitch_pad
    big = torch.full((n, ld), float("nan"), device="cuda")
    big[:, :d] = torch.from_numpy(X).cuda()
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    return big, [[big[i, offs[p]:offs[p + 1]] for p in range(len(seg))] for i in range(n)]


@pytest.mark.parametrize("impl", ["default", "ldg", "bulk", "bulk256"])
@pytest.mark.parametrize("n", [2, 3, 5, 10, 16, 20, 31, 32, 36, 48, 64, 72, 96, 100, 128])
def test_median_and_trimmed_dense_tma_path(n, impl, monkeypatch):
    """K1 fed by TMA (2-D tensor-map boxes on a dense matrix; per-row 1-D bulk copies when forced
    with FLAGG_K1_IMPL=bulk) against the LDG kernel and the oracle: bit-exact median including
    +-0 / NaN / Inf columns, ragged tails (tensor lengths not multiples of 128 columns, zero-filled
    by the TMA unit) and several tensors per client."""
    from flagg import MEDIAN, TRIMMED_MEAN, engine
    monkeypatch.delenv("FLAGG_K1_DENSE_COLS", raising=False)
    if impl == "default":
        monkeypatch.delenv("FLAGG_K1_IMPL", raising=False)
    elif impl == "bulk256":                                            # 256-column boxes at every n they serve (default: 80..110)
        if n > 110:
            pytest.skip("256-column tiles serve n <= 110")
        monkeypatch.setenv("FLAGG_K1_IMPL", "bulk")
        monkeypatch.setenv("FLAGG_K1_DENSE_COLS", "256")
    else:
        monkeypatch.setenv("FLAGG_K1_IMPL", impl)
    rng = np.random.RandomState(4000 + n)
    seg = [256, 4, 1000, 132, 120]                       # every offset a multiple of 4 floats
    d = sum(seg)
    X = rows_model(rng, n, d, byz=n // 5)
    X[:, :150] = rng.choice(EDGE, size=(n, 150))
    X[:, 150:182] = rng.choice(np.array([0.0, -0.0], dtype=np.float32), size=(n, 32))
    Xi = X.view(np.uint32)
    for j in range(182, 200):
        Xi[rng.randint(n), j] = rng.choice([0x7FC00000, 0xFFC00001, 0x7F800123])
    big, lists = dense_lists(X, seg)
    u = engine.Updates(lists)
    got = engine.colwise_select(u, MEDIAN).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X))), (n, impl)
    k = max(1, n // 10)
    if n - 2 * k >= 1:
        Xt = X.copy()
        Xt[:, :200] = rows_model(rng, n, 200)            # finite data for the tolerance comparison
        big[:, :200] = torch.from_numpy(Xt[:, :200]).cuda()
        gt = engine.colwise_select(u, TRIMMED_MEAN, k).cpu().numpy()
        assert_close_scaled(gt, oracle64.trimmed_mean64(Xt, k), oracle64.trimmed_scale(Xt, k), tol=2e-6, what=f"n={n} k={k} {impl}")
    u.close()


@pytest.mark.parametrize("gather_cols", ["256", "128"])
@pytest.mark.parametrize("n", [2, 3, 5, 10, 20, 31, 33, 64, 100, 110, 112, 128])
def test_median_and_trimmed_gather4_path(n, gather_cols, monkeypatch):
    """K1 on SEPARATELY ALLOCATED rows (one allocation per (client, tensor), 512-byte aligned: the
    reference's list-of-lists, run_experiments.py:238): the tile is fed by tile::gather4 requests through
    the view-wide tensor maps -- 256-column overlapped rows (4 KB per request) or 128-column rows.
    Covers n not a multiple of 4 (padding rows), ragged last units with <= 128 and > 128 columns left
    (the per-row fallback inside the wide kernel), tiny tensors, and must equal the LDG feed bit for bit."""
    from flagg import MEDIAN, TRIMMED_MEAN, engine
    monkeypatch.setenv("FLAGG_K1_GATHER_COLS", gather_cols)
    monkeypatch.delenv("FLAGG_K1_IMPL", raising=False)
    rng = np.random.RandomState(5000 + n)
    seg = [256 * 3 + 130, 7, 128, 256 * 2 + 100, 255, 257, 1]
    d = sum(seg)
    X = rows_model(rng, n, d, byz=n // 5)
    X[:, :150] = rng.choice(EDGE, size=(n, 150))
    X[:, 150:182] = rng.choice(np.array([0.0, -0.0], dtype=np.float32), size=(n, 32))
    Xi = X.view(np.uint32)
    for j in range(182, 200):
        Xi[rng.randint(n), j] = rng.choice([0x7FC00000, 0xFFC00001, 0x7F800123])
    ups = dev_lists(X, seg)
    assert all(t.data_ptr() % 512 == 0 for row in ups for t in row)
    u = engine.Updates(ups)
    got = engine.colwise_select(u, MEDIAN).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X))), (n, gather_cols)
    monkeypatch.setenv("FLAGG_K1_IMPL", "ldg")
    assert np.array_equal(bits(got), bits(engine.colwise_select(u, MEDIAN).cpu().numpy()))
    monkeypatch.delenv("FLAGG_K1_IMPL")
    k = max(1, n // 10)
    if n - 2 * k >= 1:
        Xt = X.copy()
        Xt[:, :200] = rows_model(rng, n, 200)
        u2 = engine.Updates(dev_lists(Xt, seg))
        gt = engine.colwise_select(u2, TRIMMED_MEAN, k).cpu().numpy()
        assert_close_scaled(gt, oracle64.trimmed_mean64(Xt, k), oracle64.trimmed_scale(Xt, k), tol=2e-6, what=f"n={n} k={k}")
        u2.close()
    u.close()


def test_gather_map_needs_512_byte_alignment_else_ldg():
    """Rows at a 16-byte offset inside their allocations cannot use the gather map (a ragged last
    unit could be read past its granule); the view falls back to the LDG feed and FLAGG_K1_IMPL=bulk
    says so instead of silently doing something else."""
    from flagg import FlaggError, MEDIAN, engine
    import os
    rng = np.random.RandomState(77)
    n, d = 20, 1000
    X = rows_model(rng, n, d)
    # allocations of different sizes: the rows are not equally spaced (that would be a dense matrix: 2-D TMA boxes)
    ups = [[torch.zeros(d + 4 + 1024 * (i % 3), device="cuda")[4:4 + d].copy_(torch.from_numpy(X[i]))] for i in range(n)]
    assert len({ups[i + 1][0].data_ptr() - ups[i][0].data_ptr() for i in range(n - 1)}) > 1
    u = engine.Updates(ups)
    assert np.array_equal(bits(engine.colwise_select(u, MEDIAN).cpu().numpy()), bits(oracle64.median_stable(X)))
    os.environ["FLAGG_K1_IMPL"] = "bulk"
    try:
        with pytest.raises(FlaggError):
            engine.colwise_select(u, MEDIAN)
    finally:
        del os.environ["FLAGG_K1_IMPL"]
    u.close()


@pytest.mark.parametrize("impl", ["tile", "radix"])
@pytest.mark.parametrize("n", [129, 130, 131, 200, 257, 1000, 2047])
def test_median_large_n_tile_and_radix(n, impl, monkeypatch):
    """n beyond the register networks: the shared-memory tile select (colselect_tile.cu, default)
    and the old re-reading radix select (FLAGG_K1_IMPL=radix), bit-exact against the
    (value, index)-ordered median incl. +-0 / NaN / Inf / duplicate columns and ragged tiles."""
    from flagg import MEDIAN
    monkeypatch.setenv("FLAGG_K1_IMPL", impl)
    rng = np.random.RandomState(n)
    d = 300
    X = rows_model(rng, n, d, byz=n // 10)
    X[:, :100] = rng.choice(EDGE, size=(n, 100))
    X.view(np.uint32)[rng.randint(n), 101] = 0x7FC00001
    got = run_select(X, MEDIAN, seg=[260, 40])
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X)))


@pytest.mark.parametrize("n,k", [(3, 1), (5, 1), (5, 2), (10, 1), (20, 2), (20, 4), (33, 9), (64, 6), (100, 10),
                                 (100, 28), (128, 12), (129, 12), (130, 64), (300, 30), (1000, 100), (1000, 499), (2047, 1)])
@pytest.mark.parametrize("impl", ["auto", "tile", "radix"])
def test_trimmed_mean_vs_oracle(n, k, impl, monkeypatch):
    from flagg import TRIMMED_MEAN
    if impl != "auto":
        if n <= 128:
            pytest.skip("FLAGG_K1_IMPL=tile|radix only matters beyond the register networks")
        monkeypatch.setenv("FLAGG_K1_IMPL", impl)
    rng = np.random.RandomState(7 * n + k)
    d = 1100 if n <= 128 else 260
    X = rows_model(rng, n, d, byz=min(k, n // 5))
    X[:k // 2 + 1] *= 100.0                                          # scaling attackers
    X[:, :64] = rng.choice(np.array([0.0, -0.0, 1.0, -1.0, 2.5, 1e-38, 1e30], dtype=np.float32), size=(n, 64))
    X.view(np.uint32)[rng.randint(n), 70] = 0x7FC00000               # one NaN: sorts last and is trimmed
    got = run_select(X, TRIMMED_MEAN, k, seg=[d - 100, 100])
    ref, kk, fb = ref_port.trimmed_mean(torch.from_numpy(X), k / n + 1e-9)
    assert kk == k and not fb
    assert_close_scaled(got, ref.numpy(), oracle64.trimmed_scale(X, k), what=f"n={n} k={k}")
    # and against the FP64 statement, tighter: one rounding of the exact mean
    t64 = oracle64.trimmed_mean64(X, k)
    assert_close_scaled(got, t64, oracle64.trimmed_scale(X, k), tol=2e-7, what="fp64")


def test_large_n_select_unaligned_views_and_all_equal_columns(monkeypatch):
    """Tile select on rows that are 4-byte aligned views (4-byte cp.async path), with columns whose
    values are all equal (no informative bit) and columns of exact duplicates around the median."""
    from flagg import MEDIAN, TRIMMED_MEAN, engine
    rng = np.random.RandomState(77)
    n, d = 300, 1001
    X = rows_model(rng, n, d, byz=30)
    X[:, 5] = 0.125
    X[:, 6] = -0.0
    X[100:220, 7] = X[100, 7]
    big = torch.from_numpy(X).cuda()                                  # pitch 1001 floats: rows 4-byte aligned only
    u = engine.Updates([[big[i]] for i in range(n)])
    monkeypatch.setenv("FLAGG_K1_IMPL", "tile")
    got = engine.colwise_select(u, MEDIAN, 0).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X)))
    gt = engine.colwise_select(u, TRIMMED_MEAN, 30).cpu().numpy()
    assert_close_scaled(gt, oracle64.trimmed_mean64(X, 30), oracle64.trimmed_scale(X, 30), tol=2e-7, what="tile trimmed")
    u.close()


def test_trimmed_mean_nan_kept_and_inf():
    from flagg import TRIMMED_MEAN
    n, k, d = 9, 1, 130
    X = np.random.RandomState(3).normal(size=(n, d)).astype(np.float32)
    X[2, 0] = np.nan; X[5, 0] = np.nan           # two NaNs, k=1: one survives the trim -> NaN
    X[1, 1] = np.inf; X[7, 1] = np.inf           # +inf kept -> +inf
    X[1, 2] = np.inf; X[3, 2] = -np.inf; X[4, 2] = np.inf; X[6, 2] = -np.inf   # +inf and -inf kept -> NaN
    got = run_select(X, TRIMMED_MEAN, k)
    ref = ref_port.trimmed_mean(torch.from_numpy(X), k / n + 1e-9)[0].numpy()
    assert np.array_equal(np.isnan(got), np.isnan(ref))
    assert np.isnan(got[0]) and np.isposinf(got[1]) and np.isnan(got[2])
    fin = np.isfinite(ref)
    np.testing.assert_allclose(got[fin], ref[fin], rtol=1e-5, atol=1e-6)


def test_select_unaligned_and_ragged_segments():
    """Rows that are 4-byte-aligned views (odd offsets into one buffer) and tiny / empty segments."""
    from flagg import MEDIAN, engine
    rng = np.random.RandomState(11)
    n, seg = 7, [1, 0, 3, 129, 127, 2, 1025]
    d = sum(seg)
    X = rng.normal(size=(n, d)).astype(np.float32)
    big = torch.zeros(n * (d + 3) + 8, device="cuda")
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    lists = []
    for i in range(n):
        base = i * (d + 3) + 1                                       # +1 float: only 4-byte aligned
        big[base:base + d] = torch.from_numpy(X[i]).cuda()
        lists.append([big[base + offs[p]:base + offs[p + 1]] for p in range(len(seg))])
    u = engine.Updates(lists)
    got = engine.colwise_select(u, MEDIAN).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle64.median_stable(X)))
    st = engine.row_stats(u, torch.from_numpy(X[0]).cuda())
    np.testing.assert_allclose(st["sumsq"], oracle64.row_sumsq64(X), rtol=1e-6)
    np.testing.assert_allclose(st["dot"], oracle64.row_dot64(X, X[0]), rtol=1e-5, atol=1e-4)
    w = rng.uniform(0.5, 2, n).astype(np.float32)
    comb = engine.row_combine(u, None, w, None, 3.0).cpu().numpy()
    assert np.array_equal(bits(comb), bits(oracle64.combine_seq32(X, w, None, 3.0)))
    u.close()


@pytest.mark.parametrize("n,d,seg", [(10, 421642, [288, 32, 18432, 64, 401408, 128, 1280, 10]),   # C1: SimpleCNN
                                     (37, 50000, [50000]), (500, 20011, [20000, 11])])
def test_row_stats_and_combine_vs_oracle(n, d, seg):
    from flagg import engine
    rng = np.random.RandomState(n)
    X = rows_model(rng, n, d, byz=2)
    g = rng.normal(0, 1e-3, d).astype(np.float32)
    u = engine.Updates(dev_lists(X, seg))
    st = engine.row_stats(u, torch.from_numpy(g).cuda())
    np.testing.assert_allclose(st["sumsq"], oracle64.row_sumsq64(X), rtol=2e-7)
    np.testing.assert_allclose(st["dot"], oracle64.row_dot64(X, g), rtol=1e-5, atol=1e-9)
    assert abs(st["g0_sumsq"] - float(np.dot(g.astype(np.float64), g.astype(np.float64)))) <= 2e-7 * st["g0_sumsq"]
    assert np.array_equal(st["linf"], np.abs(X).max(axis=1))
    rows = rng.permutation(n)[: max(1, n // 2)].astype(np.int32)
    w1 = rng.uniform(0.1, 3, len(rows)).astype(np.float32)
    w2 = rng.uniform(0.1, 3, len(rows)).astype(np.float32)
    got = engine.row_combine(u, rows, w1, w2, 7.25).cpu().numpy()
    assert np.array_equal(bits(got), bits(oracle64.combine_seq32(X, w1, w2, 7.25, rows=rows.tolist())))
    u.close()


@pytest.mark.parametrize("n,d,f", [(10, 421642, 2), (7, 1000, 1), (65, 3000, 10), (130, 2049, 20), (200, 777, 50)])
def test_gram_distances_scores_vs_fp64(n, d, f):
    from flagg import engine
    rng = np.random.RandomState(d)
    X = rows_model(rng, n, d, byz=f)
    X[1] = X[0]                                                       # colluders: exact duplicate rows
    u = engine.Updates(dev_lists(X))
    G = engine.gram(u)
    D, scores = engine.krum_scores(G, f)
    D64 = oracle64.distances64(X)
    assert D[0, 1] == 0.0 and np.all(np.diag(D) == 0) and np.array_equal(D, D.T)
    mask = D64 > 0
    assert np.max(np.abs(D[mask] - D64[mask]) / D64[mask]) <= 1e-5    # north_star: Gram-derived distances 1e-5
    s64 = oracle64.krum_scores64(D64, f)
    np.testing.assert_allclose(scores, s64, rtol=1e-5)
    assert scores[0] == scores[1]
    for mk in (1, max(2, n // 2)):
        assert np.argsort(scores)[:mk].tolist() == np.argsort(s64)[:mk].tolist() or \
            set(np.argsort(scores)[:mk].tolist()) == set(np.argsort(s64)[:mk].tolist())
    u.close()


@pytest.mark.parametrize("impl", ["tc", "ss", "ts1", "ts2", "ffma"])
@pytest.mark.parametrize("n,seg,f", [(64, [4096], 10), (128, [32], 20), (65, [3000], 10), (130, [2048, 1], 20),
                                     (200, [780, 64, 5000], 50), (257, [33, 95, 4100], 60), (600, [5000, 77, 1300], 100),
                                     (1000, [7000, 36, 3000], 100)])
def test_gram_paths_vs_fp64(n, seg, f, impl, monkeypatch):
    """K4 on the tensor cores (tcgen05 3xTF32: "tc" = the dispatcher's choice, "ss" = 128x128 tiles
    with both operands in shared memory, "ts2" = 256x256
    CTA-pair tiles, gram_tc.cu / gram_ts.cuh / gram_ts2.cuh) and on the CUDA cores (gram.cu), forced
    through FLAGG_GRAM_IMPL, against FP64: distances within 1e-5 relative (north_star), exact
    symmetry, zero distance between exact duplicates, identical Krum selection.  Ragged segment
    lengths exercise zero-padded k-blocks and row tiles beyond n."""
    from flagg import engine
    monkeypatch.setenv("FLAGG_GRAM_IMPL", impl)
    d = sum(seg)
    rng = np.random.RandomState(n + d)
    X = rows_model(rng, n, d, byz=f)
    X[1] = X[0]
    u = engine.Updates(dev_lists(X, seg))
    G = engine.gram(u)
    D, scores = engine.krum_scores(G, f)
    D64 = oracle64.distances64(X)
    assert D[0, 1] == 0.0 and np.all(np.diag(D) == 0) and np.array_equal(D, D.T)
    mask = D64 > 0
    assert np.max(np.abs(D[mask] - D64[mask]) / D64[mask]) <= 1e-5    # tolerance stated by north_star
    s64 = oracle64.krum_scores64(D64, f)
    np.testing.assert_allclose(scores, s64, rtol=1e-5)
    for mk in (1, max(2, n // 2)):
        assert set(np.argsort(scores)[:mk].tolist()) == set(np.argsort(s64)[:mk].tolist())
    u.close()


@pytest.mark.parametrize("impl", ["auto", "small", "ffma", "ts1"])
@pytest.mark.parametrize("n,seg,f", [(4, [4096], 0), (8, [777, 33], 2), (9, [5000, 1, 260], 2), (10, [288, 32, 18432, 64, 401408, 128, 1280, 10], 2),
                                     (16, [3000], 4), (17, [64, 2050], 4), (20, [130000, 36], 5), (25, [4100, 95], 6),
                                     (32, [9000, 31], 8), (33, [4096, 100], 8), (50, [3000, 36], 12)])
def test_gram_small_cohorts_vs_fp64(n, seg, f, impl, monkeypatch):
    """K4 for the reference's default cohort sizes: n <= 32 runs the register-tile kernel
    (gram_small.cu), 33 <= n < 64 the single-CTA tensor kernel; each forced kernel and the
    dispatcher's choice against FP64 -- distances within 1e-5 relative, exact symmetry, zero
    distance between duplicates, identical Krum selection.  Ragged segments (1, 31, 33, 95
    columns) exercise partial 32-column groups."""
    from flagg import engine
    if impl == "small" and n > 32:
        pytest.skip("gram_small serves n <= 32")
    if impl == "auto":
        monkeypatch.delenv("FLAGG_GRAM_IMPL", raising=False)
    else:
        monkeypatch.setenv("FLAGG_GRAM_IMPL", impl)
    d = sum(seg)
    rng = np.random.RandomState(n + d)
    X = rows_model(rng, n, d, byz=f)
    X[1] = X[0]
    u = engine.Updates(dev_lists(X, seg))
    G = engine.gram(u)
    D, scores = engine.krum_scores(G, f)
    D64 = oracle64.distances64(X)
    assert D[0, 1] == 0.0 and np.all(np.diag(D) == 0) and np.array_equal(D, D.T)
    mask = D64 > 0
    assert np.max(np.abs(D[mask] - D64[mask]) / D64[mask]) <= 1e-5    # tolerance stated by north_star
    if n - f - 2 >= 1:
        s64 = oracle64.krum_scores64(D64, f)
        np.testing.assert_allclose(scores, s64, rtol=1e-5)
        assert int(np.argmin(scores)) == int(np.argmin(s64))
    u.close()


@pytest.mark.parametrize("aligned", [True, False, "dense", "offset16"])
@pytest.mark.parametrize("n,seg,f", [(2, [300], 0), (3, [4096], 0), (4, [4096], 0), (8, [777, 33], 2), (9, [5000, 1, 260], 2),
                                     (10, [288, 32, 18432, 64, 401408, 128, 1280, 10], 2), (16, [3000], 4), (17, [64, 2050], 4),
                                     (20, [130000, 36], 5), (24, [50000], 6), (25, [4100, 95], 6), (31, [700, 3], 8), (32, [9000, 31], 8)])
@pytest.mark.parametrize("impl", ["default", "jobs", "pairs"])
def test_sqdist_small_cohorts_vs_fp64(n, seg, f, aligned, impl, monkeypatch):
    """flagg_sqdist (the reference's default cohort sizes, n <= 32): squared distances by direct
    differences, both kernels (row-block jobs: every job layout, 1 / 2 / 3 / 4 row blocks, 128-bit and
    4-byte copies; pair groups: 1-5 groups, bulk copies and the ragged-unit path), ragged tensors.
    Distances within 1e-5 of FP64 (measured ~1e-7), exact symmetry and zero diagonal, exactly 0
    between duplicate rows, Krum ranking as FP64's, repeatable bit for bit, accumulate=1 adds."""
    from flagg import engine
    from flagg._lib import check, lib
    if impl == "pairs" and not aligned:
        pytest.skip("the pair-group kernel needs 16-byte aligned rows")
    if impl != "default":
        monkeypatch.setenv("FLAGG_SQDIST_IMPL", impl)
    d = sum(seg)
    rng = np.random.RandomState(n + d)
    X = rows_model(rng, n, d, byz=f)
    if n > 2:
        X[1] = X[0]
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    if aligned is True:                                                # separately allocated tensors: tile::gather4 feed
        ups = dev_lists(X, seg)
    elif aligned == "dense":                                           # rows of one matrix per tensor: 2-D box feed
        mats = [torch.from_numpy(np.ascontiguousarray(np.pad(X[:, offs[p]:offs[p + 1]], ((0, 0), (0, (-seg[p]) % 4)))).astype(np.float32)).cuda()
                for p in range(len(seg))]
        ups = [[mats[p][i, :seg[p]] for p in range(len(seg))] for i in range(n)]
    elif aligned == "offset16":                                        # 16-byte but not 512-byte aligned rows: per-row bulk copies
        bufs = [[torch.zeros(seg[p] + 4, device="cuda") for p in range(len(seg))] for _ in range(n)]
        ups = [[bufs[i][p][4:] for p in range(len(seg))] for i in range(n)]
        for i in range(n):
            for p in range(len(seg)):
                ups[i][p].copy_(torch.from_numpy(X[i, offs[p]:offs[p + 1]]))
    else:                                                              # 4-byte aligned views into one buffer
        big = torch.zeros((n, d + 1), device="cuda")
        big[:, 1:] = torch.from_numpy(X).cuda()
        ups = [[big[i, 1 + offs[p]:1 + offs[p + 1]] for p in range(len(seg))] for i in range(n)]
    u = engine.Updates(ups)
    D2 = engine.sqdist(u)
    D, scores = engine.krum_scores(D2, f, want_scores=n - f - 2 >= 1, squared=True)
    D64 = oracle64.distances64(X)
    assert np.all(np.diag(D) == 0) and np.array_equal(D, D.T)
    assert n == 2 or D[0, 1] == 0.0
    mask = D64 > 0
    err = np.max(np.abs(D[mask] - D64[mask]) / D64[mask])
    assert err <= 1e-5, err
    assert err <= 2e-6                                                 # no cancellation: an fp32 sum of at most 1024 squares per lane
    if n - f - 2 >= 1:
        s64 = oracle64.krum_scores64(D64, f)
        np.testing.assert_allclose(scores, s64, rtol=1e-5)
        assert int(np.argmin(scores)) == int(np.argmin(s64))
    again = engine.sqdist(u)
    assert torch.equal(D2, again)
    check(lib().flagg_sqdist(u.view(), 1, again.data_ptr(), torch.cuda.current_stream().cuda_stream), "flagg_sqdist")
    assert torch.equal(again, 2.0 * D2)
    u.close()


@pytest.mark.parametrize("impl", ["jobs", "pairs"])
def test_sqdist_nonfinite_rows_follow_ieee_like_torch_norm(impl, monkeypatch):
    """A NaN client, an Inf client and a pair of Inf clients: flagg_sqdist gives NaN / inf exactly
    where the reference's torch.norm(f_i - f_j) does (krum.py:95), no special path involved."""
    from flagg import engine
    monkeypatch.setenv("FLAGG_SQDIST_IMPL", impl)
    rng = np.random.RandomState(3)
    n, d = 14, 3000
    X = rows_model(rng, n, d)
    X[2, 17] = np.nan
    X[5, 100] = np.inf
    X[9, 100] = np.inf                                                # inf - inf = NaN between rows 5 and 9
    X[11, 2999] = -np.inf
    u = engine.Updates(dev_lists(X, [1000, 2000]))
    D, _ = engine.krum_scores(engine.sqdist(u), 3, squared=True)
    with np.errstate(all="ignore"):
        ref = ref_port.krum_distances(torch.from_numpy(X))
    assert np.array_equal(np.isnan(D), np.isnan(ref)) and np.array_equal(np.isinf(D), np.isinf(ref))
    fin = np.isfinite(ref) & (ref > 0)
    assert np.max(np.abs(D[fin] - ref[fin]) / ref[fin]) <= 1e-5
    with pytest.raises(engine.FlaggError):
        engine.sqdist(engine.Updates(dev_lists(rows_model(rng, 33, 64), [64])))
    u.close()


@pytest.mark.parametrize("n", [1, 2, 3, 31])
def test_gram_small_many_tiny_tensors(n, monkeypatch):
    """Hundreds of tensors of 1..300 elements (biases, norm scales): every 128-column unit is
    partial, stages straddle tensors, and the centre row changes alignment per tensor."""
    from flagg import engine
    monkeypatch.delenv("FLAGG_GRAM_IMPL", raising=False)
    rng = np.random.RandomState(100 + n)
    seg = [int(x) for x in rng.randint(1, 301, size=257)]
    d = sum(seg)
    X = rows_model(rng, n, d, byz=0)
    u = engine.Updates(dev_lists(X, seg))
    c = torch.from_numpy(X[0] + np.float32(1e-4)).cuda()
    for center in (None, c):
        G = engine.gram(u, center=center, auto_center=False).cpu().numpy()
        Xc = X.astype(np.float64) - (0.0 if center is None else center.cpu().numpy().astype(np.float64))
        G64 = Xc @ Xc.T
        assert np.array_equal(G, G.T)
        assert np.max(np.abs(G - G64) / np.sqrt(np.outer(np.diag(G64), np.diag(G64)))) <= 1e-5
    u.close()


def test_gram_small_on_unaligned_views_and_accumulate(monkeypatch):
    """The small-cohort kernel takes 4-byte aligned row views (scalar loads) and honours
    accumulate=1 (the sharded path adds shard Grams): G(X) + G(X) == 2 G(X) bit for bit."""
    from flagg import engine
    monkeypatch.delenv("FLAGG_GRAM_IMPL", raising=False)
    rng = np.random.RandomState(5)
    n, d = 12, 5001
    X = rows_model(rng, n, d, byz=2)
    big = torch.from_numpy(X).cuda()
    u = engine.Updates([[big[i]] for i in range(n)])
    c = torch.from_numpy(np.median(X, axis=0)).cuda()
    G = engine.gram(u, center=c).cpu().numpy()
    Xc = X.astype(np.float64) - c.cpu().numpy().astype(np.float64)
    G64 = Xc @ Xc.T
    assert np.array_equal(G, G.T)
    assert np.max(np.abs(G - G64) / np.sqrt(np.outer(np.diag(G64), np.diag(G64)))) <= 1e-6
    G2 = engine.gram(u, center=c).cpu().numpy()
    assert np.array_equal(G, G2)                                       # deterministic reduction order
    from flagg._lib import check, lib
    Gd = torch.from_numpy(G).cuda()
    check(lib().flagg_gram(u.view(), c.data_ptr(), 1, Gd.data_ptr(), torch.cuda.current_stream().cuda_stream), "flagg_gram")
    assert np.array_equal(Gd.cpu().numpy(), 2.0 * G)                   # accumulate=1 adds the same sums
    monkeypatch.setenv("FLAGG_GRAM_IMPL", "small")
    big2 = torch.from_numpy(rows_model(rng, 40, 512, byz=4)).cuda()
    u40 = engine.Updates([[big2[i]] for i in range(40)])
    with pytest.raises(engine.FlaggError):                             # the forced kernel refuses n > 32 loudly
        engine.gram(u40)
    u40.close()
    u.close()


def test_gram_tc_is_default_for_large_n_and_falls_back_when_unaligned(monkeypatch):
    """n >= 64 on 16-byte aligned tensors takes the tensor-core path; rows that are 4-byte
    aligned views take the CUDA-core path; both agree to 1e-5 on every distance."""
    from flagg import engine
    monkeypatch.delenv("FLAGG_GRAM_IMPL", raising=False)
    rng = np.random.RandomState(3)
    n, d = 96, 5001
    X = rows_model(rng, n, d, byz=9)
    big = torch.from_numpy(X).cuda()                                   # row stride 5001 floats: rows 4-byte aligned only
    u_unal = engine.Updates([[big[i]] for i in range(n)])
    u_al = engine.Updates(dev_lists(X))
    c = torch.from_numpy(np.median(X[::4], axis=0)).cuda()
    Ga, Gu = engine.gram(u_al, center=c).cpu().numpy(), engine.gram(u_unal, center=c).cpu().numpy()
    Xc = X.astype(np.float64) - c.cpu().numpy().astype(np.float64)
    G64 = Xc @ Xc.T
    scale = np.sqrt(np.outer(np.diag(G64), np.diag(G64)))
    assert np.max(np.abs(Ga - G64) / scale) <= 1e-5 and np.max(np.abs(Gu - G64) / scale) <= 1e-6
    assert not np.array_equal(Ga, Gu)                                  # different arithmetic => different paths ran
    monkeypatch.setenv("FLAGG_GRAM_IMPL", "tc")
    with pytest.raises(engine.FlaggError):
        engine.gram(u_unal, center=c)
    u_al.close(); u_unal.close()


@pytest.mark.parametrize("impl,n", [("ss", 96), ("ts1", 96), ("ts1", 300), ("ts2", 96), ("ts2", 300)])
def test_gram_tensor_core_paths_on_16_byte_aligned_views(impl, n, monkeypatch):
    """Rows that are views of one matrix with a 16-byte (not 32-byte) multiple pitch: the
    tensor-core producers must take their 128-bit load path; centred Gram against FP64."""
    from flagg import engine
    monkeypatch.setenv("FLAGG_GRAM_IMPL", impl)
    rng = np.random.RandomState(11 + n)
    d = 5004                                                           # pitch 20016 B = 16 * 1251
    X = rows_model(rng, n, d, byz=n // 10)
    big = torch.from_numpy(X).cuda()
    u = engine.Updates([[big[i]] for i in range(n)])
    c = torch.from_numpy(np.median(X[::4], axis=0)).cuda()
    G = engine.gram(u, center=c).cpu().numpy()
    Xc = X.astype(np.float64) - c.cpu().numpy().astype(np.float64)
    G64 = Xc @ Xc.T
    assert np.array_equal(G, G.T)
    assert np.max(np.abs(G - G64) / np.sqrt(np.outer(np.diag(G64), np.diag(G64)))) <= 1e-5
    u.close()


def test_krum_selected_set_matches_reference_port_c1():
    """BASELINE config 1 shape: 10 clients x SimpleCNN (421,642 params), Krum f=2."""
    from src.defenses import KrumDefense
    rng = np.random.RandomState(1)
    seg = [288, 32, 18432, 64, 401408, 128, 1280, 10]
    X = rows_model(rng, 10, sum(seg), byz=2)
    for mk in (1, 5):
        d = KrumDefense({"num_malicious": 2, "multi_k": mk})
        res = d.aggregate(dev_lists(X, seg), [100] * 10)
        ref = ref_port.krum(torch.from_numpy(X), 2, mk)
        assert d.selected_clients == ref["selected"]
        np.testing.assert_allclose(d.client_scores, ref["scores"], rtol=1e-5)
        got = torch.cat([t.flatten().cpu() for t in res]).numpy()
        assert np.array_equal(bits(got), bits(ref["out"].numpy()))


def test_host_path_equals_device_path():
    from flagg import MEDIAN, TRIMMED_MEAN, engine
    rng = np.random.RandomState(5)
    n, seg = 20, [70000, 333, 200000]
    X = rows_model(rng, n, sum(seg), byz=4)
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    host = [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).pin_memory() for p in range(3)] for i in range(n)]
    ud, uh = engine.Updates(dev_lists(X, seg)), engine.Updates(host)
    for mode, k in ((MEDIAN, 0), (TRIMMED_MEAN, 2)):
        a = engine.colwise_select(ud, mode, k).cpu().numpy()
        b = engine.colwise_select(uh, mode, k)
        assert not b.is_cuda and np.array_equal(bits(a), bits(b.numpy()))
    ud.close()


def test_host_path_uniform_pitch_rows_cached_pipe_and_release():
    """Host rows that are views of ONE pinned matrix move with 2-D copies; the staging pipe is
    reused across calls (same shapes), rebuilt when n changes, and freed by flagg_host_release."""
    from flagg import MEDIAN, TRIMMED_MEAN, engine
    from flagg._lib import lib
    rng = np.random.RandomState(6)
    for n, d in ((20, 300_001), (20, 300_001), (33, 150_000)):
        X = rows_model(rng, n, d, byz=3)
        host = torch.from_numpy(X).pin_memory()
        offs = [0, 1000, 1000 + 77, d]
        uh = engine.Updates([[host[i, offs[p]:offs[p + 1]] for p in range(3)] for i in range(n)])
        ud = engine.Updates(dev_lists(X, [offs[p + 1] - offs[p] for p in range(3)]))
        for mode, k in ((MEDIAN, 0), (TRIMMED_MEAN, 3)):
            a = engine.colwise_select(ud, mode, k).cpu().numpy()
            b = engine.colwise_select(uh, mode, k)
            assert not b.is_cuda and np.array_equal(bits(a), bits(b.numpy()))
        ud.close()
    assert lib().flagg_host_release() == 0 and lib().flagg_host_release() == 0


def test_native_communicator_single_rank_and_errors():
    """flagg_comm (NCCL through the C-ABI) with one rank: sum / max all-reduce are identities and
    stream-ordered; sharded defenses accept it as 'process_group'.  The multi-rank check is
    tools/comm_check.py under torchrun (profiles/r01_comm_check_n2.txt)."""
    from flagg import Comm, FlaggError, engine
    from src.defenses import KrumDefense
    comm = Comm(1, 0, Comm.unique_id())
    a = torch.arange(1000, dtype=torch.float64, device="cuda") * 0.5
    b = torch.arange(77, dtype=torch.float32, device="cuda") - 30
    ra, rb = a.clone(), b.clone()
    comm.all_reduce(a, "sum"); comm.all_reduce(b, "max")
    torch.cuda.synchronize()
    assert torch.equal(a, ra) and torch.equal(b, rb)
    with pytest.raises(ValueError):
        comm.all_reduce(torch.zeros(4, dtype=torch.int32, device="cuda"))
    with pytest.raises(FlaggError):
        Comm(2, 5, Comm.unique_id())                                  # rank outside [0, nranks)
    rng = np.random.RandomState(9)
    X = rows_model(rng, 12, 5000, byz=2)
    ups = dev_lists(X)
    k0, k1 = KrumDefense({"num_malicious": 2, "multi_k": 3}), KrumDefense({"num_malicious": 2, "multi_k": 3, "process_group": comm})
    k0.aggregate(ups, [1] * 12); k1.aggregate(ups, [1] * 12)
    assert k0.selected_clients == k1.selected_clients and k0.client_scores == k1.client_scores
    comm.close()


def test_empty_column_shard_joins_every_step():
    """A rank whose column shard is empty (d < 1024 x ranks, flagg.sharding.shard_update_lists hands it one
    zero-length tensor per client) must run every defense to the end -- zero partial statistics into the
    all-reduces, an empty slice out -- instead of raising and leaving the other ranks in a collective.  One rank and
    a one-rank communicator are enough to walk every code path."""
    from flagg import Comm
    from src.defenses import (FLTrustDefense, GeometricMedianDefense, GradientClippingDefense, KrumDefense, MedianDefense,
                              MultiKrumDefense, NoDefense, NormBoundingDefense, TrimmedMeanDefense, DPSGDDefense)

    class Fixed(FLTrustDefense):
        def _compute_server_gradient(self, global_params):
            self.server_gradient = global_params
            return global_params

    comm = Comm(1, 0, Comm.unique_id())
    for n in (12, 40):
        ups = [[torch.zeros(0, device="cuda")] for _ in range(n)]
        ne = [3] * n
        cases = [(KrumDefense, {"num_malicious": 2, "multi_k": 1}, {}), (MultiKrumDefense, {"num_malicious": 2, "multi_k": 4}, {}),
                 (MedianDefense, {}, {}), (TrimmedMeanDefense, {"trim_ratio": 0.1}, {}), (NoDefense, {}, {}),
                 (GradientClippingDefense, {"clip_norm": 1.0}, {}), (GradientClippingDefense, {"clip_norm": 1.0, "clip_type": "linf"}, {}),
                 (NormBoundingDefense, {"max_norm": 1.0}, {}), (DPSGDDefense, {"clip_norm": 1.0, "noise_multiplier": 0.1}, {}),
                 (GeometricMedianDefense, {"max_iters": 5}, {}),
                 (Fixed, {}, {"global_params": [torch.zeros(0, device="cuda")]})]
        for cls, cfg, kw in cases:
            res = cls(dict(cfg, process_group=comm)).aggregate(ups, ne, **kw)
            assert sum(int(t.numel()) for t in res) == 0, cls.__name__
    torch.cuda.synchronize()
    comm.close()


def test_errors_surface_as_exceptions():
    from flagg import FlaggError, TRIMMED_MEAN, engine
    from src.defenses import KrumDefense, get_defense
    u = engine.Updates(dev_lists(np.zeros((4, 16), dtype=np.float32)))
    with pytest.raises(FlaggError):
        engine.colwise_select(u, TRIMMED_MEAN, 2)                     # n - 2k < 1
    with pytest.raises(ValueError):
        KrumDefense({"num_malicious": 2}).aggregate(dev_lists(np.zeros((3, 8), dtype=np.float32)), [1] * 3)
    with pytest.raises(ValueError):
        get_defense("nope", {})
    u.close()


# ---- full-size properties (BASELINE config 3: 100 x 25M, median, 10 GB) -----------------
def test_full_size_median_properties():
    from flagg import MEDIAN, engine
    n, d = 100, 25_000_000
    g = torch.Generator(device="cuda").manual_seed(1234)
    theta = torch.randn(d, device="cuda", generator=g) * 0.05
    X = torch.empty((n, d), device="cuda")
    for i in range(n):
        X[i] = theta + torch.randn(d, device="cuda", generator=g) * (1e-2 if i < 10 else 1e-3)
    del theta
    u = engine.Updates([[X[i]] for i in range(n)])
    out = engine.colwise_select(u, MEDIAN)
    # (1) sampled columns against the oracle, bit-exact
    cols = torch.randint(0, d, (50_000,), device="cuda", generator=g)
    sub = X[:, cols].cpu().numpy()
    assert np.array_equal(bits(out[cols].cpu().numpy()), bits(oracle64.median_stable(sub)))
    # (2) rank property over ALL columns: #less <= r and #less_or_equal >= r+1
    r = (n - 1) // 2
    lt = torch.zeros(d, dtype=torch.int16, device="cuda")
    le = torch.zeros(d, dtype=torch.int16, device="cuda")
    for i in range(n):
        lt += (X[i] < out)
        le += (X[i] <= out)
    assert bool((lt <= r).all()) and bool((le >= r + 1).all())
    # (3) permutation invariance: client order does not change a tie-free median
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(1)).tolist()
    u2 = engine.Updates([[X[i]] for i in perm])
    out2 = engine.colwise_select(u2, MEDIAN)
    assert torch.equal(out.view(torch.int32), out2.view(torch.int32))
    u.close(); u2.close()
