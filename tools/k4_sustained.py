#!/usr/bin/env python
"""Per-launch time of the C4-shard Gram over 30 back-to-back launches, with SM clock samples (GPU box only)."""
import os, sys, subprocess
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from flagg import engine
n, d = 1000, 3_125_000
g = torch.Generator(device="cuda").manual_seed(1)
theta = torch.randn(d, device="cuda", generator=g) * 0.05
X = torch.randn((n, d), device="cuda", generator=g) * 1e-3 + theta
u = engine.Updates([[X[i]] for i in range(n)])
for name, c in (("centre = theta", theta.clone()), ("centre = robust_center", engine.robust_center(u))):
    engine.gram(u, center=c); torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(31)]
    ev[0].record()
    for i in range(30):
        engine.gram(u, center=c); ev[i + 1].record()
    torch.cuda.synchronize()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(30)]
    clk = subprocess.run(["nvidia-smi", "--query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active", "--format=csv,noheader"], capture_output=True, text=True).stdout.strip()
    print(name, "first 3:", [round(t, 2) for t in ts[:3]], "last 3:", [round(t, 2) for t in ts[-3:]], "min %.2f max %.2f" % (min(ts), max(ts)), "| after:", clk, flush=True)
