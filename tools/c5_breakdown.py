#!/usr/bin/env python
"""C5 round breakdown on one GPU: K2 with / without g0, K3 over all rows / kept rows, the whole clipping and FLTrust rounds."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import ctypes as C
import numpy as np
import torch
from flagg import engine
from flagg._lib import check, lib

n, d = 500, 11_689_512
X = torch.randn((n, d), device="cuda") * 1e-3
g0 = torch.randn(d, device="cuda") * 1e-3
X[50:] += 0.7 * g0
ups = [[X[i]] for i in range(n)]
u = engine.Updates(ups)
st = torch.cuda.current_stream().cuda_stream
buf = torch.zeros(2 * n + 1, dtype=torch.float64, device="cuda")
linf = torch.zeros(n, dtype=torch.float32, device="cuda")
out = torch.empty(d, device="cuda")
p = buf.data_ptr()


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


L = lib()
t = timed(lambda: check(L.flagg_row_reduce(u.view(), g0.data_ptr(), 0, p, p + 8 * n, linf.data_ptr(), p + 16 * n, st), "k2"))
print(f"K2 with g0 (sumsq, dot, linf)      {t:7.3f} ms  {4.0 * n * d / t / 1e6:7.0f} GB/s")
t = timed(lambda: check(L.flagg_row_reduce(u.view(), None, 0, p, None, linf.data_ptr(), None, st), "k2"))
print(f"K2 without g0 (sumsq, linf)        {t:7.3f} ms  {4.0 * n * d / t / 1e6:7.0f} GB/s")
w = np.ones(n, dtype=np.float32)
t = timed(lambda: engine.row_combine(u, None, w, w, float(n)))
print(f"K3 all 500 rows (host weights)     {t:7.3f} ms  {4.0 * n * d / t / 1e6:7.0f} GB/s")
rows = list(range(50, n))
t = timed(lambda: engine.row_combine(u, rows, w[50:], w[50:], float(n)))
print(f"K3 450 rows                        {t:7.3f} ms  {4.0 * 450 * d / t / 1e6:7.0f} GB/s")
ne = [50 + i % 7 for i in range(n)]
t = timed(lambda: engine.weighted_round(u, engine.W_CLIP, ne, p0=0.0035))
print(f"weighted_round CLIP                {t:7.3f} ms")
t = timed(lambda: engine.weighted_round(u, engine.W_FLTRUST, None, g0))
print(f"weighted_round FLTRUST             {t:7.3f} ms")
