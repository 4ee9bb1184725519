#!/usr/bin/env python
"""A few flagg_sqdist launches at (n, d) for ncu:   sqdist_once.py n [d] [jobs|pairs]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from flagg import engine

n = int(sys.argv[1])
d = int(sys.argv[2]) if len(sys.argv) > 2 else 11_689_512
if len(sys.argv) > 3:
    os.environ["FLAGG_SQDIST_IMPL"] = sys.argv[3]
ups = [[torch.randn(d, device="cuda") * 1e-3 + 0.05] for _ in range(n)]
u = engine.Updates(ups)
for _ in range(6):
    D2 = engine.sqdist(u)
torch.cuda.synchronize()
print(n, d, float(D2[0, 1]))
