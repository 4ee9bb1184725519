#!/usr/bin/env python
"""K1 at (n, d) on the three input layouts: rows of ONE dense matrix (2-D TMA boxes), separately
allocated 512-byte-aligned rows (the reference's list-of-lists: tile::gather4, four clients per TMA
request) and separately allocated rows at a 16-byte offset (LDG feed).  Kernel time through the
C-ABI on a prebuilt view and the same call through MedianDefense.aggregate.
    k1_layouts.py [n] [d] [median|trimmed]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import json

import torch
from flagg import MEDIAN, TRIMMED_MEAN, engine
from src.defenses import MedianDefense, TrimmedMeanDefense

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100
d = int(sys.argv[2]) if len(sys.argv) > 2 else 25_000_000
mode = TRIMMED_MEAN if len(sys.argv) > 3 and sys.argv[3] == "trimmed" else MEDIAN
k = max(1, n // 10) if mode == TRIMMED_MEAN else 0
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6534.8


def timed(fn, steps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


X = torch.randn((n, d), device="cuda") * 1e-3 + 0.05
layouts = {
    "dense": lambda: [[X[i]] for i in range(n)],
    "separate_512B_aligned": lambda: [[torch.empty(d, device="cuda").copy_(X[i])] for i in range(n)],
    "separate_16B_offset": lambda: [[torch.empty(d + 4, device="cuda")[4:].copy_(X[i])] for i in range(n)],
}
ref = None
for name, make in layouts.items():
    ups = make()
    u = engine.Updates(ups)
    out = engine.colwise_select(u, mode, k)
    if ref is None:
        ref = out.clone()
    same = bool(torch.equal(out.view(torch.int32), ref.view(torch.int32)))
    t_kernel = timed(lambda: engine.colwise_select(u, mode, k))
    defense = MedianDefense({}) if mode == MEDIAN else TrimmedMeanDefense({"trim_ratio": 0.1})
    t_api = timed(lambda: defense.aggregate(ups, [1] * n))
    alg = 4.0 * d * (n + 1)
    print(f"n={n} d={d} {name:24s} kernel {t_kernel:7.3f} ms = {alg / t_kernel / 1e6:7.0f} GB/s = {alg / t_kernel / 1e6 / peak:5.3f} of HBM | "
          f"plugin API {t_api:7.3f} ms = {alg / t_api / 1e6 / peak:5.3f} | identical to dense: {same}", flush=True)
    u.close()
    del ups, u
