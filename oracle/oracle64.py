"""Oracle, level 2: library-independent numpy restatement (FP64 where it matters).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).

Two jobs:
  * an independent statement of the *selection* semantics the CUDA kernels must
    reproduce bit-exactly -- coordinate-wise median under torch's total order
    (value ascending, client index ascending, +0 == -0, any NaN -> first NaN;
    torch.median(dim=0) as reached from trimmed_mean.py:163) -- pinned against
    torch in tests/test_oracle_golden.py;
  * FP64 truth for norms / distances / weighted sums at large d, where the
    reference's own fp32 ``torch.norm`` is biased low and
    "within 1e-5 of the reference" is not reachable by a correct kernel.
"""
from __future__ import annotations

import numpy as np


def median_stable(X: np.ndarray) -> np.ndarray:
    """Rank-(n-1)//2 element of every column under (value, client index) order.

    Follows what ``torch.median(stacked, dim=0)[0]`` (trimmed_mean.py:163) returns
    on CPU: NaN anywhere in the column -> the first NaN (payload kept); otherwise
    the lower median, ties (including -0 vs +0) broken by client index.
    """
    X = np.ascontiguousarray(X, dtype=np.float32)
    n, d = X.shape
    r = (n - 1) // 2
    order = np.argsort(X, axis=0, kind="stable")          # -0 == +0 compares equal; NaN last
    out = np.take_along_axis(X, order[r:r + 1], axis=0)[0].copy()
    nan_mask = np.isnan(X)
    has_nan = nan_mask.any(axis=0)
    if has_nan.any():
        first = nan_mask.argmax(axis=0)
        cols = np.nonzero(has_nan)[0]
        out[cols] = X[first[cols], cols]
    return out


def trimmed_mean64(X: np.ndarray, k: int) -> np.ndarray:
    """Mean of sorted ranks [k, n-k) per column, accumulated in FP64 (NaN sorts last,
    as torch.sort does at trimmed_mean.py:82)."""
    n = X.shape[0]
    s = np.sort(np.asarray(X, dtype=np.float64), axis=0)
    return s[k:n - k].mean(axis=0)


def trimmed_scale(X: np.ndarray, k: int) -> np.ndarray:
    """mean_j |v_j| over the kept summands -- the scale of the tolerance metric
    |got-ref| <= 1e-5 * (|ref| + scale)  ('Tolerance definition')."""
    n = X.shape[0]
    s = np.sort(np.asarray(X, dtype=np.float64), axis=0)
    return np.abs(s[k:n - k]).mean(axis=0)


def row_sumsq64(X: np.ndarray) -> np.ndarray:
    X = np.asarray(X)
    return np.einsum("ij,ij->i", X, X, dtype=np.float64)


def row_dot64(X: np.ndarray, g: np.ndarray) -> np.ndarray:
    return np.asarray(X, dtype=np.float32) @ np.asarray(g, dtype=np.float64)


def distances64(X: np.ndarray, block: int = 1 << 20) -> np.ndarray:
    """Pairwise un-squared L2 distances (krum.py:95 computes these in fp32) in FP64,
    by direct differences (no Gram trick), column-blocked to bound memory."""
    X = np.asarray(X, dtype=np.float32)
    n, d = X.shape
    D2 = np.zeros((n, n))
    for c0 in range(0, d, block):
        B = X[:, c0:c0 + block].astype(np.float64)
        for i in range(n):
            diff = B[i + 1:] - B[i]
            D2[i, i + 1:] += np.einsum("ij,ij->i", diff, diff)
    D2 = D2 + D2.T
    return np.sqrt(D2)


def krum_scores64(D: np.ndarray, f: int) -> np.ndarray:
    """krum.py:126-129 on an FP64 distance matrix."""
    n = D.shape[0]
    m = n - f - 2
    return np.sort(D, axis=1)[:, 1:m + 1].sum(axis=1)


def weighted_combine64(X: np.ndarray, w: np.ndarray, divisor: float) -> np.ndarray:
    return (np.asarray(w, dtype=np.float64) @ np.asarray(X, dtype=np.float64)) / divisor


def combine_seq32(X: np.ndarray, w1, w2, divisor, rows=None) -> np.ndarray:
    """Strict left-to-right fp32 restatement of the reference's ``sum()`` aggregation
    (base_defense.py:91-95; fltrust.py:264-268; differential_privacy.py:265-269):
    out = (((0 + w2_0*(w1_0*x_0)) + w2_1*(w1_1*x_1)) + ...) / divisor with every
    multiply and add rounded to fp32 separately.  ``w1``/``w2`` None = stage absent."""
    X = np.asarray(X, dtype=np.float32)
    rows = range(X.shape[0]) if rows is None else rows
    acc = np.zeros(X.shape[1], dtype=np.float32)
    for t, i in enumerate(rows):
        term = X[i]
        if w1 is not None:
            term = term * np.float32(w1[t])
        if w2 is not None:
            term = term * np.float32(w2[t])
        acc = acc + term
    return acc / np.float32(divisor)
