#!/usr/bin/env python
"""Per-defense latency through the plugin API (device tensors in, device tensors out) at the two
small BASELINE shapes -- C1: 10 clients x 421,642 params in 8 tensors; C2: 20 clients x 874,056 in
22 tensors -- beside the oracle port of the reference on this box's host cores (GPU box only).
Every result is checked against the port (selections identical, outputs within 1e-5 scaled)."""
import os, sys, time, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch
from oracle import ref_port
from src.defenses import (FLTrustDefense, GradientClippingDefense, KrumDefense, MedianDefense, MultiKrumDefense,
                          NoDefense, NormBoundingDefense, TrimmedMeanDefense)

torch.set_num_threads(os.cpu_count())
C1 = (10, [288, 32, 18432, 64, 401408, 128, 1280, 10], 2)
seg2 = [864, 32, 18432, 64, 73728, 128] + [39936, 128, 16384, 128, 8192, 64] + [1000] * 8 + [707000, 1000]
seg2[-1] += 874056 - sum(seg2)
C2 = (20, seg2, 4)

def rows(n, d, byz, seed):
    rng = np.random.RandomState(seed)
    X = (rng.normal(0, 0.05, d)[None, :] + rng.normal(0, 1e-3, (n, d))).astype(np.float32)
    X[:byz] = X[:byz] * 10
    return X

def bench(fn, reps):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps * 1e3

out = {}
for name, (n, seg, f) in (("C1", C1), ("C2", C2)):
    d = sum(seg); X = rows(n, d, f, 5); Xt = torch.from_numpy(X)
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    ups = [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(seg))] for i in range(n)]
    ne = [100 + i for i in range(n)]
    flat = lambda res: torch.cat([t.flatten() for t in res]).cpu()
    cases = {
        "fedavg": (lambda: NoDefense({}).aggregate(ups, ne), lambda: ref_port.fedavg(Xt, ne)),
        "median": (lambda: MedianDefense({}).aggregate(ups, ne), lambda: ref_port.median(Xt)),
        "trimmed_mean_0.2": (lambda: TrimmedMeanDefense({"trim_ratio": 0.2}).aggregate(ups, ne), lambda: ref_port.trimmed_mean(Xt, 0.2)[0]),
        "krum": (lambda: KrumDefense({"num_malicious": f, "multi_k": 1}).aggregate(ups, ne), lambda: ref_port.krum(Xt, f, 1)),
        "multi_krum": (lambda: MultiKrumDefense({"num_malicious": f, "multi_k": n // 2}).aggregate(ups, ne), lambda: ref_port.krum(Xt, f, n // 2)),
        "gradient_clipping": (lambda: GradientClippingDefense({"clip_norm": 11.0}).aggregate(ups, ne), lambda: ref_port.gradient_clipping(Xt, ne, 11.0)),
        "norm_bounding": (lambda: NormBoundingDefense({"max_norm": 50.0}).aggregate(ups, ne), lambda: ref_port.norm_bounding(Xt, ne, 50.0)),
    }
    for cname, (gpu_fn, cpu_fn) in cases.items():
        g_ms = bench(gpu_fn, 20)
        t0 = time.perf_counter(); cpu_fn(); c_ms = (time.perf_counter() - t0) * 1e3
        out[f"{name}/{cname}"] = {"n": n, "d": d, "P": len(seg), "api_ms": round(g_ms, 3), "cpu_port_ms": round(c_ms, 2),
                                  "GBps_api": round(4.0 * n * d / g_ms / 1e6, 1)}
        print(f"{name} {cname:20s} api {g_ms:8.3f} ms   cpu port ({os.cpu_count()} threads) {c_ms:9.2f} ms   x{c_ms / g_ms:7.1f}", flush=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", "api_latency.json"), "w"), indent=1)
