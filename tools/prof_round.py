import os, sys, time, cProfile, pstats
ROOT="/root/repo"
sys.path[:0]=[ROOT, os.path.join(ROOT,"multimodal-fl-security_b200")]
import torch
from src.defenses import FLTrustDefense, GradientClippingDefense, MultiKrumDefense
class Fixed(FLTrustDefense):
    def _compute_server_gradient(self, g):
        self.server_gradient = g
        return g
n,d=500,8192
X=torch.randn((n,d),device="cuda")*1e-3
g0=torch.randn(d,device="cuda")*1e-3
X[50:]+=0.7*g0
ups=[[X[i]] for i in range(n)]
ne=[50+i%7 for i in range(n)]
fl=Fixed({}); gc=GradientClippingDefense({"clip_norm":0.1})
for name,fn in (("fltrust",lambda: fl.aggregate(ups,ne,global_params=[g0])),("clip",lambda: gc.aggregate(ups,ne))):
    for _ in range(5): fn()
    torch.cuda.synchronize(); t0=time.perf_counter()
    for _ in range(50): fn()
    torch.cuda.synchronize(); print(name, (time.perf_counter()-t0)/50*1e3, "ms per call at n=500, d=8192")
pr=cProfile.Profile(); pr.enable()
for _ in range(50): fl.aggregate(ups,ne,global_params=[g0])
pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(18)
