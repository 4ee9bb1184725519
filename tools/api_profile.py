#!/usr/bin/env python
"""cProfile of one small-shape aggregate() call (GPU box only): where the fixed latency goes."""
import os, sys, cProfile, pstats
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch
from src.defenses import KrumDefense, MedianDefense
n, seg = 10, [288, 32, 18432, 64, 401408, 128, 1280, 10]
X = (np.random.RandomState(0).normal(0, 0.05, (n, sum(seg)))).astype(np.float32)
offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
ups = [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(seg))] for i in range(n)]
for cls, cfg in ((MedianDefense, {}), (KrumDefense, {"num_malicious": 2, "multi_k": 1})):
    d = cls(cfg)
    for _ in range(3): d.aggregate(ups, [1] * n)
    torch.cuda.synchronize()
    pr = cProfile.Profile(); pr.enable()
    for _ in range(50): d.aggregate(ups, [1] * n)
    torch.cuda.synchronize(); pr.disable()
    print("=====", cls.__name__)
    pstats.Stats(pr).sort_stats("tottime").print_stats(12)
