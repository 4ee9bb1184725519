#!/usr/bin/env python
"""K1 timing for n beyond the register-network sizes (GPU box only): median / trimmed mean over
n clients x d columns, GB/s of client updates and fraction of the measured HBM copy rate."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from flagg import MEDIAN, TRIMMED_MEAN, engine
peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
for n, d in ((128, 6_250_000), (200, 6_250_000), (256, 6_250_000), (500, 3_125_000), (1000, 3_125_000), (2000, 1_000_000)):
    X = torch.randn((n, d), device="cuda") * 1e-3
    u = engine.Updates([[X[i]] for i in range(n)])
    for mode, k, name in ((MEDIAN, 0, "median"), (TRIMMED_MEAN, n // 10, "trimmed")):
        engine.colwise_select(u, mode, k); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3): engine.colwise_select(u, mode, k)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 3
        gbs = 4.0 * (n + 1) * d / ms / 1e6
        print(f"n={n:5d} d={d:8d} {name:8s} {ms:8.2f} ms  {gbs:7.0f} GB/s  {100*gbs/peak:5.1f}% of measured HBM", flush=True)
    u.close(); del X, u
