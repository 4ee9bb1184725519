#!/usr/bin/env python
"""cProfile of one small-shape aggregate() call (GPU box only): where the fixed latency goes."""
import os, sys, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch
from src.defenses import KrumDefense, MedianDefense
n, seg = 10, [288, 32, 18432, 64, 401408, 128, 1280, 10]
if "C2" in sys.argv:   # 20 clients x 874,056 params in 22 tensors
    seg = [864, 32, 18432, 64, 73728, 128] + [39936, 128, 16384, 128, 8192, 64] + [1000] * 8 + [707000, 1000]
    seg[-1] += 874056 - sum(seg)
    n = 20
X = (np.random.RandomState(0).normal(0, 0.05, (n, sum(seg)))).astype(np.float32)
offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
ups = [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(seg))] for i in range(n)]
for cls, cfg in ((MedianDefense, {}), (KrumDefense, {"num_malicious": 2 if n == 10 else 4, "multi_k": 1})):
    d = cls(cfg)
    for _ in range(3): d.aggregate(ups, [1] * n)
    torch.cuda.synchronize()
    pr = cProfile.Profile(); pr.enable()
    for _ in range(50): d.aggregate(ups, [1] * n)
    torch.cuda.synchronize(); pr.disable()
    print("=====", cls.__name__)
    pstats.Stats(pr).sort_stats("tottime").print_stats(12)
