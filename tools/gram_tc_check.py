#!/usr/bin/env python
"""Dev check of the tcgen05 Gram path against FP64 numpy + timing (GPU box only)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch
from flagg import engine

def rows_model(rng, n, d, byz):
    theta = rng.normal(0, 0.05, d)
    X = theta[None, :] + rng.normal(0, 1e-3, (n, d))
    X[:byz] = theta[None, :] + rng.normal(0, 1e-2, (byz, d))
    return X.astype(np.float32)

def check(n, segs, impl):
    os.environ["FLAGG_GRAM_IMPL"] = impl
    d = sum(segs)
    rng = np.random.RandomState(n * 7 + d)
    X = rows_model(rng, n, d, max(1, n // 10))
    offs = np.concatenate([[0], np.cumsum(segs)]).astype(int)
    Xd = torch.from_numpy(X).cuda()
    ups = [[Xd[i, offs[p]:offs[p + 1]].clone() for p in range(len(segs))] for i in range(n)]
    u = engine.Updates(ups)
    c = engine.robust_center(u)
    G = engine.gram(u, center=c).cpu().numpy()
    Xc = X.astype(np.float64) - c.cpu().numpy().astype(np.float64)[None, :]
    G64 = Xc @ Xc.T
    sc = np.sqrt(np.outer(np.diag(G64), np.diag(G64)))
    eg = np.max(np.abs(G - G64) / sc)
    D2 = np.diag(G)[:, None] + np.diag(G)[None, :] - 2 * G
    D264 = np.diag(G64)[:, None] + np.diag(G64)[None, :] - 2 * G64
    m = ~np.eye(n, dtype=bool)
    ed = np.max(np.abs(np.sqrt(np.maximum(D2[m], 0)) - np.sqrt(D264[m])) / np.sqrt(D264[m]))
    print(f"impl={impl:4s} n={n:5d} d={d:9d} P={len(segs):2d}  max|dG|/sqrt(GiiGjj)={eg:.2e}  max rel dist err={ed:.2e}  sym={np.array_equal(G, G.T)}", flush=True)
    u.close()
    return eg, ed

def timing(n, d, impl, reps=3):
    os.environ["FLAGG_GRAM_IMPL"] = impl
    g = torch.Generator(device="cuda").manual_seed(1)
    theta = torch.randn(d, device="cuda", generator=g) * 0.05
    X = torch.randn((n, d), device="cuda", generator=g) * 1e-3 + theta
    u = engine.Updates([[X[i]] for i in range(n)])
    c = theta.clone()
    engine.gram(u, center=c); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        engine.gram(u, center=c)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"impl={impl:4s} n={n} d={d}: {ms:.2f} ms  alg {n*(n+1)*d/ms/1e9:.1f} TFLOP/s  hw(3xTF32 upper tiles) "
          f"{3*2*(((n+127)//128)*(((n+127)//128)+1)//2)*128*128*d/ms/1e9:.1f} TFLOP/s  {4*n*d/ms/1e6:.0f} GB/s", flush=True)
    u.close()

def tf32_peak():
    torch.backends.cuda.matmul.allow_tf32 = True
    a = torch.randn(8192, 8192, device="cuda"); b = torch.randn(8192, 8192, device="cuda")
    best = 1e9
    for _ in range(12):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); torch.matmul(a, b); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"cuBLAS TF32 8192^3 best of 12: {2*8192**3/best/1e9:.1f} TFLOP/s", flush=True)

if __name__ == "__main__":
    bad = 0
    if "--only-time" in sys.argv:
        timing(int(os.environ.get("K4_N", 1000)), int(os.environ.get("K4_D", 3_125_000)), sys.argv[-1] if sys.argv[-1] in ("tc", "ss", "ts1", "ts2", "small", "ffma") else "tc", reps=2)
        sys.exit(0)
    if "--peak" in sys.argv:
        tf32_peak()
    IMPLS = [a for a in sys.argv[1:] if a in ("tc", "ss", "ts1", "ts2", "ffma")] or ["tc", "ss", "ts2", "ffma"]
    for n, segs in [(64, [4096]), (257, [4096 + 32]), (600, [5000, 77, 130000]), (128, [32]), (128, [256 * 40]), (65, [3000]), (130, [2049 - 1, 1]), (200, [777 + 3, 64, 5000]),
                    (300, [100000]), (1000, [70000, 36, 30000])]:
        for impl in IMPLS:
            eg, ed = check(n, segs, impl)
            bad += (ed > 1e-5) or not np.isfinite(ed)
    if "--time" in sys.argv:
        timing(1000, 3_125_000, "tc")
        timing(1000, 3_125_000, "ss")
        timing(500, 3_125_000, "tc")
        timing(128, 3_125_000, "tc")
        timing(1000, 3_125_000, "ffma", reps=1)
        timing(100, 25_000_000 // 8, "tc")
        timing(500, 1_461_188, "tc")
    print("BAD" if bad else "ALL OK")
    sys.exit(1 if bad else 0)
