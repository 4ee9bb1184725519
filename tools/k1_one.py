#!/usr/bin/env python
"""One K1 launch at (n, d) for ncu: k1_one.py n d [median|trimmed]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from flagg import MEDIAN, TRIMMED_MEAN, engine
n, d = int(sys.argv[1]), int(sys.argv[2])
mode = TRIMMED_MEAN if len(sys.argv) > 3 and sys.argv[3] == "trimmed" else MEDIAN
X = torch.randn((n, d), device="cuda") * 1e-3 + 0.05
u = engine.Updates([[X[i]] for i in range(n)])
for _ in range(2):
    engine.colwise_select(u, mode, n // 10 if mode == TRIMMED_MEAN else 0)
torch.cuda.synchronize()
