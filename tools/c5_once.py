#!/usr/bin/env python
"""FLTrust at C5 (500 x 11,689,512) through FLTrustDefense.aggregate, a few rounds: the K2 / K3 launches ncu captures."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
from src.defenses import FLTrustDefense


class Fixed(FLTrustDefense):
    def _compute_server_gradient(self, global_params):
        self.server_gradient = global_params
        return global_params


n, d = 500, 11_689_512
g0 = torch.randn(d, device="cuda") * 1e-3
X = torch.randn((n, d), device="cuda") * 1e-3
X[50:] += 0.7 * g0
ups = [[X[i]] for i in range(n)]
fl = Fixed({})
for _ in range(4):
    fl.aggregate(ups, [1] * n, global_params=[g0])
torch.cuda.synchronize()
print("ok", sum(1 for t in fl.trust_scores if t > 0))
