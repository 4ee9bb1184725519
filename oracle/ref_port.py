"""Oracle, level 1: torch-CPU restatement of the reference defenses in matrix form.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Every function takes the
client updates as one fp32 CPU matrix ``X[n, d]`` (row i = client i's tensors
flattened and concatenated) and follows the reference's own sequence of torch /
numpy calls, because several results depend on library internals (median tie
rule, ``torch.norm`` accumulation, ``mean`` summation order, ``argsort`` tie
order).  Element-wise defenses act per tensor in the reference; acting on the
concatenation is the same arithmetic.  Citations are file:line under
/root/reference/src/defenses/.
"""
from __future__ import annotations

import numpy as np
import torch


def _f32_scalar(v) -> torch.Tensor:
    """A Python int/float scalar as torch applies it to an fp32 tensor (cast to fp32)."""
    return torch.tensor(float(v), dtype=torch.float64).to(torch.float32)


def _left_to_right(terms):
    """Python ``sum()`` over tensors: ``0 + t0 + t1 + ...`` strictly in order
    (base_defense.py:91-94, krum.py:189, fltrust.py:264-267)."""
    acc = None
    for t in terms:
        acc = (torch.zeros_like(t) + t) if acc is None else acc + t
    return acc


# ----------------------------------------------------------------------------
# FedAvg  (base_defense.py:80-97)
# ----------------------------------------------------------------------------
def fedavg(X: torch.Tensor, num_examples) -> torch.Tensor:
    total = sum(num_examples)
    acc = _left_to_right(int(num_examples[i]) * X[i] for i in range(X.shape[0]))
    return acc / total


# ----------------------------------------------------------------------------
# Median / trimmed mean  (trimmed_mean.py:48-103, 141-166)
# ----------------------------------------------------------------------------
def median(X: torch.Tensor) -> torch.Tensor:
    """trimmed_mean.py:160-163 -- stack().float() then torch.median(dim=0)[0]."""
    return torch.median(X.float(), dim=0)[0]


def trim_count(n: int, trim_ratio: float) -> int:
    """trimmed_mean.py:66 -- note the double rounding of n*ratio before int()."""
    return max(1, int(n * trim_ratio))


def trimmed_mean(X: torch.Tensor, trim_ratio: float):
    """trimmed_mean.py:63-90.  Returns (result, k, fell_back_to_median)."""
    n = X.shape[0]
    k = trim_count(n, trim_ratio)
    if n - 2 * k < 1:                       # :69-72 fallback
        return median(X), k, True
    sorted_vals, _ = torch.sort(X.float(), dim=0)      # :82
    return sorted_vals[k:n - k].mean(dim=0), k, False  # :85-88


# ----------------------------------------------------------------------------
# Krum / Multi-Krum  (krum.py:73-192)
# ----------------------------------------------------------------------------
def krum_distances(X: torch.Tensor) -> np.ndarray:
    """krum.py:86-99 -- n(n-1)/2 fp32 torch.norm calls into a float64 matrix."""
    n = X.shape[0]
    D = np.zeros((n, n))
    for i in range(n):
        for j in range(i + 1, n):
            D[i, j] = D[j, i] = torch.norm(X[i] - X[j]).item()
    return D


def krum_scores(D: np.ndarray, f: int) -> list:
    """krum.py:122-131,160-169 -- sum of the n-f-2 smallest non-self distances."""
    n = D.shape[0]
    m = n - f - 2
    return [np.sum(np.sort(D[i].copy())[1:m + 1]) for i in range(n)]


def krum_check(n: int, f: int) -> None:
    if n < 2 * f + 3:                       # krum.py:153-157
        raise ValueError(f"Krum requires n >= 2f + 3. Got n={n}, f={f}.")


def krum_select(scores, multi_k: int):
    """krum.py:174-176 -- np.argsort default (introsort; tie order is numpy's)."""
    order = np.argsort(scores)
    return order[:multi_k].tolist(), order[multi_k:].tolist()


def krum(X: torch.Tensor, f: int, multi_k: int):
    """krum.py:149-192.  Returns dict(out, selected, rejected, scores, D)."""
    n = X.shape[0]
    krum_check(n, f)
    D = krum_distances(X)
    scores = krum_scores(D, f)
    sel, rej = krum_select(scores, multi_k)
    if multi_k == 1:
        out = X[sel[0]]                      # :178-181 (the client's own update)
    else:
        out = _left_to_right(X[i] for i in sel) / multi_k   # :187-190
    return dict(out=out, selected=sel, rejected=rej, scores=scores, D=D)


# ----------------------------------------------------------------------------
# FLTrust  (fltrust.py:152-270) -- g0 (the server gradient) is an input here;
# producing it is model training (fltrust.py:93-150), outside the hot path.
# ----------------------------------------------------------------------------
def fltrust(X: torch.Tensor, g0: torch.Tensor):
    n = X.shape[0]
    g0 = g0.float()
    trust = []
    for i in range(n):                                   # :170-186
        dot = torch.dot(X[i], g0)
        cn, sn = torch.norm(X[i]), torch.norm(g0)
        if cn < 1e-10 or sn < 1e-10:
            trust.append(0.0)
        else:
            trust.append(max(0.0, (dot / (cn * sn)).item()))
    normalized = []
    for i in range(n):                                   # :205-216
        un, rn = torch.norm(X[i]), torch.norm(g0)
        normalized.append(X[i] if un < 1e-10 else X[i] * (rn / un))
    total = sum(trust)                                   # :256
    if total < 1e-10:                                    # :258-260
        return dict(out=g0, trust=trust, fallback=True)
    acc = _left_to_right(trust[i] * normalized[i] for i in range(n))   # :264-267
    return dict(out=acc / total, trust=trust, fallback=False)


# ----------------------------------------------------------------------------
# Clipping family  (differential_privacy.py)
# ----------------------------------------------------------------------------
def _weighted_mean(rows, num_examples):
    total = sum(num_examples)
    acc = _left_to_right(int(num_examples[i]) * rows[i] for i in range(len(rows)))
    return acc / total


def gradient_clipping(X: torch.Tensor, num_examples, clip_norm: float, clip_type: str = "l2"):
    """differential_privacy.py:223-271 -- scale is a Python float here."""
    n = X.shape[0]
    norms, clipped, count = [], [], 0
    for i in range(n):
        if clip_type == "linf":
            nv = torch.max(torch.abs(X[i])).item()       # :230
        else:
            nv = torch.norm(X[i]).item()                 # :228,232
        norms.append(nv)
        if nv > clip_norm:                               # :242-245
            clipped.append(X[i] * (clip_norm / nv))
            count += 1
        else:
            clipped.append(X[i])
    return dict(out=_weighted_mean(clipped, num_examples), norms=norms, clipped_count=count)


def dp_sgd_clip_mean(X: torch.Tensor, num_examples, clip_norm: float):
    """differential_privacy.py:87-95,139-155 -- the pre-noise aggregate.  Here the
    scale ``clip_norm / norm`` is a 0-dim fp32 tensor (:91), not a Python float."""
    n = X.shape[0]
    norms, clipped = [], []
    for i in range(n):
        nrm = torch.norm(X[i])
        norms.append(nrm.item())
        clipped.append(X[i] * (clip_norm / nrm) if nrm > clip_norm else X[i])
    return dict(out=_weighted_mean(clipped, num_examples), norms=norms)


def norm_bounding(X: torch.Tensor, num_examples, max_norm: float, min_norm: float = 0.0):
    """differential_privacy.py:305-334."""
    n = X.shape[0]
    keep, rejected = [], []
    for i in range(n):
        nv = torch.norm(X[i]).item()                     # :311
        (keep if min_norm <= nv <= max_norm else rejected).append(i)
    use = keep if keep else list(range(n))               # :319-322
    out = _weighted_mean([X[i] for i in use], [num_examples[i] for i in use])
    return dict(out=out, rejected=rejected)


# ----------------------------------------------------------------------------
# Geometric median (Weiszfeld)  (trimmed_mean.py:225-265)
# ----------------------------------------------------------------------------
def geometric_median(X: torch.Tensor, max_iters: int = 100, tolerance: float = 1e-5):
    cur = torch.median(X, dim=0)[0]                      # :240
    iters = max_iters
    for it in range(max_iters):
        dist = torch.clamp(torch.norm(X - cur, dim=1), min=1e-10)   # :245-248
        w = 1.0 / dist
        new = (w.unsqueeze(1) * X).sum(dim=0) / w.sum()             # :251-252
        change = torch.norm(new - cur)
        cur = new
        if change < tolerance:                                      # :258-260
            iters = it + 1
            break
    return dict(out=cur, num_iters=iters)


# ----------------------------------------------------------------------------
# Model-poisoning attacks  (/root/reference/src/attacks/model_poisoning.py)
# ----------------------------------------------------------------------------
def attack_model_replacement(local: torch.Tensor, glob: torch.Tensor, num_clients: int, scale_factor: float = 10.0,
                             num_malicious: int = 1, boost_factor=None) -> torch.Tensor:
    """model_poisoning.py:63-78 -- old + (new - old) * scale, scale a Python float."""
    scale = boost_factor if boost_factor is not None else (num_clients / num_malicious) * scale_factor
    delta = local - glob
    return glob + delta * scale


def attack_scaling(local: torch.Tensor, scale: float = 100.0) -> torch.Tensor:
    return local * scale                                 # :236


def attack_ipm(local: torch.Tensor, epsilon: float = 0.1, benign_mean=None) -> torch.Tensor:
    if benign_mean is None:
        return -local                                    # :271-273
    return -epsilon * torch.sign(benign_mean)            # :279-281


def attack_adaptive_krum(benign: torch.Tensor, direction: torch.Tensor, perturbation_scale: float = 0.1):
    """:125-138 (centre = Python sum() of the benign rows / count) and :172-184 with a
    target_direction.  Returns (poisoned, centre)."""
    center = _left_to_right(benign[i] for i in range(benign.shape[0])) / benign.shape[0]
    return center + direction * perturbation_scale, center
