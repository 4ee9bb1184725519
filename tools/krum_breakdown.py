#!/usr/bin/env python
"""Where a sharded Multi-Krum step's time goes at the C4 shard (GPU box only)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch
from flagg import engine
n, d = 1000, 3_125_000
g = torch.Generator(device="cuda").manual_seed(1)
theta = torch.randn(d, device="cuda", generator=g) * 0.05
X = torch.randn((n, d), device="cuda", generator=g) * 1e-3 + theta
u = engine.Updates([[X[i]] for i in range(n)])
def timed(name, fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): r = fn()
    torch.cuda.synchronize()
    print(f"{name:34s} {(time.perf_counter() - t0) / reps * 1e3:8.3f} ms (wall, incl. host)", flush=True)
    return r
c = timed("robust_center", lambda: engine.robust_center(u))
G = timed("gram(center given)", lambda: engine.gram(u, center=c))
timed("gram(auto centre)", lambda: engine.gram(u))
timed("krum_scores(want_D=False)", lambda: engine.krum_scores(G, 100, want_D=False))
timed("krum_scores(want_D=True)", lambda: engine.krum_scores(G, 100, want_D=True))
rows = np.arange(900, dtype=np.int32); w = np.ones(900, dtype=np.float32)
timed("row_combine 900 rows", lambda: engine.row_combine(u, rows, w, None, 900.0))
