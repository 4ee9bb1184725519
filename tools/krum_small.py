#!/usr/bin/env python
"""Small-cohort Krum (n <= 32): flagg_sqdist kernel time, the centred-Gram path it replaces, and the
whole KrumDefense.aggregate call, at d = 11,689,512 (C2) by default.   krum_small.py [d] [n ...]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from flagg import engine
from src.defenses import KrumDefense

d = int(sys.argv[1]) if len(sys.argv) > 1 else 11_689_512
ns = [int(a) for a in sys.argv[2:]] or [4, 8, 10, 16, 20, 24, 32]


def timed(fn, steps=20, warm=5):
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


for n in ns:
    X = torch.randn((n, d), device="cuda") * 1e-3 + 0.05
    # LAYOUT=dense: rows of one matrix (2-D box feed); default: separately allocated rows (the reference's layout: gather4 feed)
    ups = [[X[i]] for i in range(n)] if os.environ.get("LAYOUT") == "dense" else [[torch.empty(d, device="cuda").copy_(X[i])] for i in range(n)]
    u = engine.Updates(ups)
    t_impl = {}
    for impl in ("jobs", "pairs"):
        os.environ["FLAGG_SQDIST_IMPL"] = impl
        try:
            engine.sqdist(u)
            t_impl[impl] = timed(lambda: engine.sqdist(u))
        except engine.FlaggError:              # the pair-group kernel serves 2..29 clients
            t_impl[impl] = float("nan")
    del os.environ["FLAGG_SQDIST_IMPL"]
    t_sq = timed(lambda: engine.sqdist(u))
    t_gram = timed(lambda: engine.gram(u))
    f = max(0, (n - 3) // 4)
    k = KrumDefense({"num_malicious": f, "multi_k": 1})
    t_api = timed(lambda: k.aggregate(ups, [1] * n)) if n >= 2 * f + 3 else float("nan")
    gb = 4.0 * n * d / 1e6
    print(f"n={n:2d} d={d} sqdist {t_sq:6.3f} ms = {gb / t_sq:6.0f} GB/s (jobs {t_impl['jobs']:6.3f}, pairs {t_impl['pairs']:6.3f}) | centre + Gram {t_gram:6.3f} ms | "
          f"KrumDefense.aggregate {t_api:6.3f} ms = {gb / t_api:6.0f} GB/s", flush=True)
    u.close()
    del X, ups, u
