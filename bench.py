#!/usr/bin/env python
"""bench.py -- aggregated client-update GB/s of the robust-aggregation hot path on B200.

    python bench.py --gpus N --steps K --warmup W            (this repo's CUDA path)
    python bench.py --impl reference --gpus N --steps K ...  (the reference's CPU path, timed
                                                              on the host cores: oracle port)

Headline (BASELINE.json north_star / config 3): coordinate-wise median of a
100-client x 25,000,000-parameter fp32 update matrix (10 GB) per GPU.  One "step" is one
aggregation of that matrix.  GB/s = 4*n*d bytes of client updates aggregated per second;
with N GPUs the parameter dimension is sharded (each rank owns its own 25M-column shard,
weak scaling) and the median needs no communication.  One JSON line on stdout (rank 0).

Beside the headline the line carries
  roofline       the dominant kernel against the measured HBM rate
  plugin_device  the same median through MedianDefense.aggregate on SEPARATELY ALLOCATED device
                 rows (the run_experiments.py:238 layout: pointer-table view, no dense matrix)
  plugin_device_resnet50  the simulator's real call: 100 clients x the 161 parameter tensors of the reference's
                 ResNet-50, every tensor its own allocation (16,100 tensor handles per aggregate() call)
  e2e            the same median through the plugin call with pinned HOST tensors (Flower path)
  e2e_pageable   Krum / FLTrust through the plugin call on pageable, separately allocated host
                 tensors (the torch.tensor(ndarray) layout of robust_server.py:99-104)
  strong         BASELINE configs 4 and 5 as STRONG scaling: the same 1000 x 25M Multi-Krum round
                 and the same 500 x 11.69M FLTrust / clipping round column-sharded over the N ranks,
                 timed through the plugin classes (all-reduce, selection and the K3 mean inside)
  defenses       the other named shapes (C2) through the plugin classes, each with a CPU baseline
  sharded_parity (N > 1) sharded selection == unsharded selection on every rank, checked before timing
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "multimodal-fl-security_b200")
for _p in (ROOT, PKG):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402
import torch  # noqa: E402

METRIC = "aggregated client-update GB/s (coordinate-wise median, 100 clients x 25M params per GPU)"
N_CLIENTS, D_PARAMS = 100, 25_000_000
WORKLOAD = "C3: 100 clients x 25,000,000 params per GPU, coordinate-wise median"
CONFIG = {"workload": WORKLOAD, "n_clients": N_CLIENTS, "n_params_per_gpu": D_PARAMS,
          "sharding": "parameter dimension over the GPUs, no collective for the median",
          "l2_policy": "inputs (10 GB) larger than L2 (126 MB); no flush needed"}
REF_BUDGET_S = 400.0        # the reference arm keeps its timed steps under about seven minutes (25 steps of <= 16 s)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def bf16_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            j = json.load(fh)
            return float(j.get("bf16_tflops", 0.0)), float(j.get("bf16_tflops_sustained", 0.0))
    return 0.0, 0.0


def tf32_peak_tflops():
    """Dense TF32 GEMM rate of this GPU (library GEMM, used only as a roofline denominator):
    best of 10 (burst) and the mean of 2 s back to back (sustained), with the SM clock sampled."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    a = torch.randn(8192, 8192, device="cuda")
    b = torch.randn(8192, 8192, device="cuda")
    best = float("inf")
    for _ in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        torch.matmul(a, b)
        e1.record()
        torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    reps = max(10, int(2000.0 / best))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        torch.matmul(a, b)
    e1.record()
    torch.cuda.synchronize()
    sustained = e0.elapsed_time(e1) / reps
    torch.backends.cuda.matmul.allow_tf32 = old
    del a, b
    flop = 2.0 * 8192 ** 3
    return flop / best / 1e9, flop / sustained / 1e9


def ncu_traffic(kernel_key):
    """dram bytes per launch from the committed ncu capture, if one exists."""
    p = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if os.path.exists(p):
        with open(p) as fh:
            return json.load(fh).get(kernel_key)
    return None


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.samples, self.proc, self.index, self.t0, self.t1 = [], None, index, None, None

    def start(self):
        """Started early (nvidia-smi needs ~0.2 s before its first line); mark() brackets the timed region."""
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "20", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append((time.perf_counter(), [x.strip() for x in line.split(",")]))

    def mark(self):
        if self.t0 is None:
            self.t0 = time.perf_counter()
        else:
            self.t1 = time.perf_counter()

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.05)
        self.proc.terminate()
        ok = [(t, r) for t, r in self.samples if len(r) == 6 and r[0].isdigit()]
        t0, t1 = self.t0 or 0.0, self.t1 or float("inf")
        rows = [r for t, r in ok if t0 <= t <= t1 + 0.03]          # samples taken while the warm-up + timed steps ran
        window = "warm-up + timed steps"
        if not rows and ok:                                         # a 40 ms region can fall between two samples: nearest ones
            rows = [r for _, r in sorted(ok, key=lambda tr: abs(tr[0] - t1))[:3]]
            window = "nearest to the timed steps"
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({nm for r in rows for nm, v in zip(names, r[2:]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median([int(r[0]) for r in rows])), "sm_max_mhz": int(rows[0][1]),
                "reasons": reasons, "samples": len(rows), "window": window}


def synth_rows(n, d, n_byz, seed, device="cuda", theta_sigma=0.05, spread=0.0):
    """Rows as run_experiments.py:238 produces them: full weights theta + delta_i.  The first
    n_byz rows are Byzantine: the reference's own ModelReplacementAttack (model_poisoning.py:48-80,
    boost_factor = 10: theta + (x - theta) * 10) applied in place on the resident rows through the
    attack mirror (flagg_segments_transform), not a hand-rolled formula."""
    from src.attacks import ModelReplacementAttack
    g = torch.Generator(device=device).manual_seed(seed)
    theta = torch.randn(d, device=device, generator=g) * theta_sigma
    sig = 1e-3 * (1.0 + spread * (2 * torch.rand(n, generator=torch.Generator().manual_seed(seed)) - 1))
    X = torch.empty((n, d), dtype=torch.float32, device=device)
    for i in range(n):
        X[i] = torch.randn(d, device=device, generator=g).mul_(float(sig[i])).add_(theta)
    attack = ModelReplacementAttack({"boost_factor": 10.0})
    for i in range(n_byz):
        attack.poison_update([X[i]], [theta], n, out=[X[i]])
    return X


def time_steps(fn, steps, warmup, dist_on):
    """W warm-up calls, then exactly K calls between CUDA events on the launching stream,
    bracketed by barrier + synchronize; returns max-over-ranks seconds."""
    import torch.distributed as dist
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], dtype=torch.float64, device="cuda")
    if dist_on:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


# ---------------------------------------------------------------------------------------------
# CPU legs: the reference's own algorithm on the host cores (oracle port; test infrastructure
# used here only as the thing timed BESIDE the product, never inside it)
# ---------------------------------------------------------------------------------------------
def _cpu_rows(n, cols, seed=1234, n_byz=None, theta_sigma=0.05):
    g = torch.Generator().manual_seed(seed)
    theta = torch.randn(cols, generator=g) * theta_sigma
    n_byz = n // 10 if n_byz is None else n_byz
    return [[theta + torch.randn(cols, generator=g) * (1e-2 if i < n_byz else 1e-3)] for i in range(n)], theta


def cpu_reference_median(steps, warmup, n=N_CLIENTS, cols=D_PARAMS, budget_s=None):
    """The reference's CPU path for the headline workload (torch.stack + torch.median(dim=0),
    trimmed_mean.py:158-164) through the oracle port, at the FULL 100 x 25M size.  With a time
    budget the step counts are cut so the run still ends in minutes; the cut is returned."""
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    rows, _ = _cpu_rows(n, cols)

    def step():
        stacked = torch.stack([u[0].float() for u in rows])      # the reference's packing copy is part of its path
        return ref_port.median(stacked)
    t0 = time.perf_counter()
    step()                                                       # first warm-up, also sizes the rest of the run
    first = time.perf_counter() - t0
    if budget_s is not None:
        afford = max(2, int(budget_s / max(first, 1e-3)))
        warmup = min(warmup, max(1, afford // 5))
        steps = min(steps, max(1, afford - warmup))
    for _ in range(max(0, warmup - 1)):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return 4.0 * n * cols / dt / 1e9, dt, torch.get_num_threads(), steps, max(1, warmup)


def cpu_trimmed_mean_c2():
    """TrimmedMeanDefense at C2 full size (20 x 11,689,512, k = 2): stack + sort + slice + mean."""
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    n, d = 20, 11_689_512
    rows, _ = _cpu_rows(n, d, seed=2, n_byz=4)

    def step():
        return ref_port.trimmed_mean(torch.stack([u[0].float() for u in rows]), 0.1)
    step()
    t0 = time.perf_counter()
    step()
    dt = time.perf_counter() - t0
    return {"value": 4.0 * n * d / dt / 1e9, "unit": "GB/s", "ms": dt * 1e3, "cores": torch.get_num_threads(), "kind": "port",
            "sample": "full size 20 x 11,689,512, 1 step after 1 warm-up (torch.stack + sort + slice + mean)"}


def cpu_krum_extrapolated(n_full, d_full, n_sub=64, d_meas=3_125_000, f_frac=0.1):
    """KrumDefense on n_sub of the n_full clients and d_meas of the d_full columns; the reference's
    cost is the n(n-1)/2 loop of torch.norm(f_i - f_j) (krum.py:92-97), linear in the row length, so
    the time is scaled by the pair count and the column count and LABELLED extrapolated."""
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    rows, _ = _cpu_rows(n_sub, d_meas, seed=3, n_byz=max(1, int(n_sub * f_frac)))
    X = torch.stack([u[0] for u in rows])
    t0 = time.perf_counter()
    ref_port.krum(X, max(1, int(n_sub * f_frac)), n_sub // 2)
    dt = time.perf_counter() - t0
    scale = (n_full * (n_full - 1)) / float(n_sub * (n_sub - 1)) * (d_full / float(d_meas))
    est = dt * scale
    return {"value": 4.0 * n_full * d_full / est / 1e9, "unit": "GB/s", "ms": est * 1e3, "cores": torch.get_num_threads(),
            "kind": "port", "extrapolated": True,
            "sample": f"{n_sub} of {n_full} clients x {d_meas} of {d_full} columns measured once ({dt:.2f} s), "
                      f"scaled by pair count x column count = x{scale:.0f}"}


def cpu_krum_full(n, d, f, multi_k):
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    rows, _ = _cpu_rows(n, d, seed=4, n_byz=f)
    X = torch.stack([u[0] for u in rows])
    t0 = time.perf_counter()
    ref_port.krum(X, f, multi_k)
    dt = time.perf_counter() - t0
    return {"value": 4.0 * n * d / dt / 1e9, "unit": "GB/s", "ms": dt * 1e3, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"full size {n} x {d}, one call (flatten + n(n-1)/2 torch.norm + argsort + mean)"}


def cpu_fltrust_sample(n, d, cols):
    """FLTrustDefense trust scores + aggregate on a column sample of C5 (time scales with columns)."""
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    rows, _ = _cpu_rows(n, cols, seed=5, n_byz=n // 10, theta_sigma=0.0)
    X = torch.stack([u[0] for u in rows])
    g0 = torch.randn(cols, generator=torch.Generator().manual_seed(6)) * 1e-3
    X[n // 10:] += 0.7 * g0
    t0 = time.perf_counter()
    ref_port.fltrust(X, g0)
    dt = time.perf_counter() - t0
    est = dt * d / cols
    return {"value": 4.0 * n * d / est / 1e9, "unit": "GB/s", "ms": est * 1e3, "cores": torch.get_num_threads(), "kind": "port",
            "extrapolated": True,
            "sample": f"{n} clients x {cols} of {d} columns measured once ({dt:.2f} s), scaled by the column count"}


# ---------------------------------------------------------------------------------------------
# GPU legs beside the headline
# ---------------------------------------------------------------------------------------------
def separate_rows(X):
    """One allocation per client (run_experiments.py:238: ``[p.data.clone() ...]``): rows at
    unrelated addresses, so flagg_view_create cannot describe them as one dense matrix."""
    return [[torch.empty(X.shape[1], dtype=torch.float32, device=X.device).copy_(X[i])] for i in range(X.shape[0])]


def resnet50_param_sizes(num_classes=200):
    """Parameter tensor sizes of the reference's CUB200CNN (torchvision ResNet-50 + a 200-way head,
    src/models/cub200_cnn.py:26-29) in model.parameters() order: 161 tensors, 23,917,832 parameters."""
    sizes = [64 * 3 * 7 * 7, 64, 64]
    inplanes = 64
    for planes, blocks in ((64, 3), (128, 4), (256, 6), (512, 3)):
        for b in range(blocks):
            sizes += [planes * inplanes, planes, planes, planes * planes * 9, planes, planes, planes * 4 * planes, planes * 4, planes * 4]
            if b == 0:
                sizes += [planes * 4 * inplanes, planes * 4, planes * 4]
            inplanes = planes * 4
    return sizes + [num_classes * 2048, num_classes]


def plugin_device_resnet50(steps):
    """The simulator's actual call (run_experiments.py:236-244): 100 clients, each update a list of the model's 161
    parameter tensors, every tensor its own allocation (``p.data.clone()``) -- 16,100 tensor handles per call."""
    from src.defenses import MedianDefense
    sizes = resnet50_param_sizes()
    n, d = N_CLIENTS, sum(sizes)
    g = torch.Generator(device="cuda").manual_seed(7)
    theta = [torch.randn(s, device="cuda", generator=g) * 0.05 for s in sizes]
    ups = [[t + torch.randn(t.numel(), device="cuda", generator=g) * (1e-2 if i < 10 else 1e-3) for t in theta] for i in range(n)]
    defense = MedianDefense({})
    t = time_steps(lambda: defense.aggregate(ups, [1] * n), steps, 3, False) / steps
    hbm, _ = peaks()
    del ups
    return {"ms": t * 1e3, "GBps": 4.0 * n * d / t / 1e9, "frac": 4.0 * d * (n + 1) / t / 1e9 / hbm, "steps": steps,
            "n_clients": n, "n_params": d, "tensors_per_client": len(sizes),
            "api": "src.defenses.MedianDefense.aggregate(100 clients x 161 separately allocated ResNet-50 parameter tensors)"}


def plugin_device_median(X, steps, dist_on, world):
    """MedianDefense.aggregate on separately allocated device rows: Updates construction, pointer-table
    upload, the per-row feed of K1, result split -- everything a caller of the plugin API pays."""
    from src.defenses import MedianDefense
    n, d = X.shape
    ups = separate_rows(X)
    defense = MedianDefense({})
    t = time_steps(lambda: defense.aggregate(ups, [1] * n), steps, 3, dist_on) / steps
    hbm, _ = peaks()
    alg = 4.0 * d * (n + 1)
    del ups
    return {"ms": t * 1e3, "GBps": world * 4.0 * n * d / t / 1e9, "frac": alg / t / 1e9 / hbm, "steps": steps,
            "api": "src.defenses.MedianDefense.aggregate(separately allocated device rows)",
            "layout": "n separate device allocations (run_experiments.py:238)"}


def e2e_median(X, out_dev, steps, dist_on, world):
    """Pinned HOST tensors in, host tensors out, through MedianDefense.aggregate (Flower path,
    robust_server.py:99-125): H2D of all n rows and D2H of the result inside the timed region."""
    import torch.distributed as dist
    from src.defenses import MedianDefense
    n, d = X.shape
    host = torch.empty((n, d), dtype=torch.float32, pin_memory=True)
    host.copy_(X)
    torch.cuda.synchronize()
    if dist_on:
        dist.barrier()
    t0 = time.perf_counter()                                    # plain pinned H2D rate with every rank copying at once
    X.copy_(host, non_blocking=True)
    torch.cuda.synchronize()
    pcie = torch.tensor([4.0 * n * d / (time.perf_counter() - t0) / 1e9], dtype=torch.float64, device="cuda")
    pcie_all = [pcie.clone() for _ in range(world)] if dist_on else [pcie]
    if dist_on:
        dist.all_gather(pcie_all, pcie)
    host_lists = [[host[i]] for i in range(n)]
    defense = MedianDefense({})
    defense.aggregate(host_lists, [1] * n)
    if dist_on:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(steps):
        res = defense.aggregate(host_lists, [1] * n)
    torch.cuda.synchronize()
    te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
    if dist_on:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    same = bool(torch.equal(res[0].view(torch.int32), out_dev.cpu().view(torch.int32)))
    del host, host_lists, res
    return {"value": world * 4.0 * n * d * steps / float(te.item()) / 1e9, "unit": "GB/s",
            "h2d_bytes_per_step": 4 * n * d, "d2h_bytes_per_step": 4 * d, "steps": steps,
            "api": "src.defenses.MedianDefense.aggregate(pinned host tensors)", "matches_device_result": same,
            "pinned_h2d_memcpy_GBps_per_gpu_concurrent": [float(p.item()) for p in pcie_all]}


def e2e_pageable(steps=3):
    """Two-pass defenses on the Flower layout: every client's tensors are separate PAGEABLE host
    allocations (robust_server.py:99-104, ``torch.tensor(arr)``); upload (worker threads gather into
    pinned bounce buffers, 8 MB copies) + kernels + D2H of the result are all inside the timed region."""
    from src.defenses import FLTrustDefense, KrumDefense

    class Fixed(FLTrustDefense):
        def _compute_server_gradient(self, global_params):
            self.server_gradient = global_params
            return global_params

    n, seg = 20, [11_000_000, 689_512]
    d = sum(seg)
    g = torch.Generator().manual_seed(77)
    theta = torch.randn(d, generator=g) * 0.05
    ups = []
    for i in range(n):
        row = theta + torch.randn(d, generator=g) * (1e-2 if i < 4 else 1e-3)
        ups.append([row[:seg[0]].clone(), row[seg[0]:].clone()])
    g0 = [torch.randn(s, generator=g) * 1e-3 for s in seg]
    out = {}
    for name, fn in (("krum_20x11.69M_f4_m10", lambda: KrumDefense({"num_malicious": 4, "multi_k": 10}).aggregate(ups, [1] * n)),
                     ("fltrust_20x11.69M", lambda: Fixed({}).aggregate(ups, [1] * n, global_params=g0))):
        fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / steps
        out[name] = {"value": 4.0 * n * d / dt / 1e9, "unit": "GB/s", "ms": dt * 1e3, "steps": steps,
                     "h2d_bytes_per_step": 4 * n * d, "d2h_bytes_per_step": 4 * d,
                     "layout": "pageable host tensors, one allocation per (client, tensor)"}
    return out


def sharded_parity_check(comm, rank, world):
    """Before any timing at N > 1: the same seeded cohort on every rank, aggregated unsharded and
    column-sharded through the communicator; selections must be identical, on every rank."""
    import torch.distributed as dist
    from flagg.sharding import shard_update_lists
    from src.defenses import FLTrustDefense, MultiKrumDefense

    class Fixed(FLTrustDefense):
        def _compute_server_gradient(self, global_params):
            self.server_gradient = global_params
            return global_params

    ok = True
    for n, d, f, mk in ((20, 400_000, 4, 10), (200, 120_000, 20, 100)):
        X = synth_rows(n, d, f, 4242, spread=0.3)
        ups = [[X[i]] for i in range(n)]
        a = MultiKrumDefense({"num_malicious": f, "multi_k": mk})
        full = a.aggregate(ups, [1] * n)[0]
        mine, (lo, hi) = shard_update_lists(ups, rank, world)
        b = MultiKrumDefense({"num_malicious": f, "multi_k": mk, "process_group": comm})
        part = torch.cat([t.reshape(-1) for t in b.aggregate(mine, [1] * n)])
        ok &= a.selected_clients == b.selected_clients and a.rejected_clients == b.rejected_clients
        ok &= bool(torch.equal(part.view(torch.int32), full[lo:hi].contiguous().view(torch.int32)))
        g0 = [X[n - 1] - X[n - 2]]
        ta, tb = Fixed({}), Fixed({"process_group": comm})
        ta.aggregate(ups, [1] * n, global_params=g0)
        tb.aggregate(mine, [1] * n, global_params=g0)
        ok &= bool(np.allclose(ta.trust_scores, tb.trust_scores, rtol=0, atol=1e-6))
        ok &= ta.detect_malicious(None, None) == tb.detect_malicious(None, None)
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(flag.item())


def c2_defenses(dist_on, with_cpu):
    """BASELINE config 2 through the plugin classes on separately allocated rows: trimmed mean
    (r = 0.1 -> k = 2) and Krum (f = 4, the reference's default cohort regime) at 20 x 11,689,512."""
    from src.defenses import KrumDefense, TrimmedMeanDefense
    hbm, _ = peaks()
    n, d = 20, 11_689_512
    X = synth_rows(n, d, 4, 99, spread=0.3)
    ups = separate_rows(X)
    del X
    out = {}
    tm = TrimmedMeanDefense({"trim_ratio": 0.1})
    t = time_steps(lambda: tm.aggregate(ups, [1] * n), 10, 3, dist_on) / 10
    out["trimmed_mean_C2_20x11.69M_k2"] = {"n": n, "d": d, "ms": t * 1e3, "GBps": 4.0 * n * d / t / 1e9, "steps": 10,
                                           "roofline_frac_hbm": 4.0 * d * (n + 1) / t / 1e9 / hbm,
                                           "api": "TrimmedMeanDefense.aggregate(separately allocated device rows)"}
    kd = KrumDefense({"num_malicious": 4, "multi_k": 1})
    t = time_steps(lambda: kd.aggregate(ups, [1] * n), 10, 3, dist_on) / 10
    out["krum_C2_20x11.69M_f4"] = {"n": n, "d": d, "ms": t * 1e3, "GBps": 4.0 * n * d / t / 1e9, "steps": 10,
                                   "roofline_frac_hbm": 4.0 * n * d / t / 1e9 / hbm,
                                   "api": "KrumDefense.aggregate(separately allocated device rows): squared distances by direct differences (flagg_sqdist) + scores + argsort"}
    del ups
    if with_cpu:
        out["trimmed_mean_C2_20x11.69M_k2"]["cpu_baseline"] = cpu_trimmed_mean_c2()
        out["krum_C2_20x11.69M_f4"]["cpu_baseline"] = cpu_krum_full(20, 11_689_512, 4, 1)
    return out


def _bcast(val, dist_on):
    """rank 0's value on every rank."""
    import torch.distributed as dist
    if dist_on:
        box = [val]
        dist.broadcast_object_list(box, src=0)
        return box[0]
    return val


def _ingest(make_rows, n, d, rank, world, dist_on, comm):
    """The round's update matrix as the sharded path meets it: ONE matrix, generated on rank 0 (the process
    that received the clients' updates), column-sharded over the ranks by flagg_comm_scatter_columns
    (ncclSend / ncclRecv per row and peer over NVLink).  Returns (this rank's shard as update lists, rank 0's
    full lists or None, (lo, hi), scatter seconds or None)."""
    import torch.distributed as dist
    from flagg.sharding import shard_bounds
    torch.cuda.empty_cache()
    full = None
    if rank == 0:
        X = make_rows()
        full = [[X[i]] for i in range(n)]
    if world == 1:
        return full, full, (0, d), None
    # NCCL builds its point-to-point connections on first use (~0.6 s): a small scatter first, then the timed one
    w = torch.zeros((8, 4096 * world), device="cuda")
    comm.scatter_columns([[w[i]] for i in range(8)] if rank == 0 else None, 8, [4096 * world], root=0)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    mine, (lo, hi) = comm.scatter_columns(full, n, [d], root=0)      # the root keeps views of its own rows
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 1e3], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    assert (lo, hi) == shard_bounds(d, world)[rank]
    return mine, full, (lo, hi), float(t.item())


def strong_scaling(rank, world, dist_on, comm, steps, with_cpu):
    """BASELINE configs 4 and 5 as strong scaling: the SAME round -- one matrix generated on rank 0 and
    column-sharded over the N ranks (ingest, reported, outside the timed aggregation) -- timed through the
    plugin classes with ``process_group=comm``; at N > 1 rank 0 then runs the whole round alone on the same
    matrix, which gives the line its own N = 1 time (efficiency = t1 / (N t_N)) and checks that the sharded
    run selected exactly what the unsharded one selects."""
    import torch.distributed as dist
    from flagg import engine
    from src.defenses import FLTrustDefense, GradientClippingDefense, MultiKrumDefense

    class Fixed(FLTrustDefense):                 # g0 injected: the root-dataset SGD step is training, not aggregation
        def _compute_server_gradient(self, global_params):
            self.server_gradient = global_params
            return global_params

    hbm, _ = peaks()
    out = {}
    # ---------------- C4: 1000 x 25,000,000, Multi-Krum f = 100, multi_k = 900 ----------------
    n, d, f, mk = 1000, 25_000_000, 100, 900
    ne = [1] * n
    mine, full, (lo, hi), t_scatter = _ingest(lambda: synth_rows(n, d, f, 31, spread=0.3), n, d, rank, world, dist_on, comm)
    defense = MultiKrumDefense({"num_malicious": f, "multi_k": mk, "process_group": comm})
    t = time_steps(lambda: defense.aggregate(mine, ne), steps, 3, dist_on) / steps
    u = engine.Updates(mine)
    tg = time_steps(lambda: engine.gram(u), 3, 1, False) / 3          # centre + this rank's Gram kernel, no collective
    u.close()
    del u
    tg_t = torch.tensor([tg], dtype=torch.float64, device="cuda")
    if dist_on:
        dist.all_reduce(tg_t, op=dist.ReduceOp.MAX)
        tg = float(tg_t.item())
    sel_ok = set(defense.rejected_clients[-f:]) == set(range(f))
    ar_us = None
    if comm is not None:
        G = torch.zeros((n, n), dtype=torch.float64, device="cuda")
        ar_us = time_steps(lambda: comm.all_reduce(G), 20, 3, True) / 20 * 1e6
        del G
    tf32_burst, tf32_sust = tf32_peak_tflops()
    bf16_burst, bf16_sust = bf16_peak()
    flops = float(n) * (n + 1) * (hi - lo)
    c4 = {"workload": "C4: 1000 clients x 25,000,000 params, Multi-Krum f=100 multi_k=900, one round",
          "api": "MultiKrumDefense.aggregate(process_group=flagg.Comm): centre + Gram + all-reduce + scores + argsort + mean of 900 rows",
          "n_gpus": world, "cols_per_gpu": hi - lo, "steps": steps, "ms": t * 1e3, "GBps": 4.0 * n * d / t / 1e9,
          "clients_x_params_per_s": n * d / t, "gram_ms_local": tg * 1e3, "allreduce_8MB_f64_us": ar_us,
          "byzantine_rows_ranked_last": sel_ok,
          "roofline": {"bound": "tensor", "kernel": "gram_ts2_kernel (tcgen05 cta_group::2 kind::tf32, 3 products per k-step)",
                       "achieved": flops / tg / 1e12, "unit": "TFLOP/s", "algorithmic_flops_per_launch": flops,
                       "peak_tf32_over_3_burst": tf32_burst / 3, "peak_tf32_over_3_sustained": tf32_sust / 3,
                       "peak_bf16_over_6_burst": bf16_burst / 6, "peak_bf16_over_6_sustained": bf16_sust / 6,
                       "frac": flops / tg / 1e12 / (tf32_burst / 3),
                       "frac_vs_bf16_over_6": (flops / tg / 1e12 / (bf16_burst / 6)) if bf16_burst else None,
                       "tf32_dense_measured_burst": tf32_burst, "tf32_dense_measured_sustained": tf32_sust}}
    if world > 1:
        c4["ingest_scatter"] = {"ms": t_scatter * 1e3, "GBps": 4.0 * n * (d - d // world) / t_scatter / 1e9,
                                "what": "rank 0 -> peers, [n, d/N] column shards, ncclSend/ncclRecv per (row, peer) over NVLink; outside the timed round"}
        sel = list(defense.selected_clients)
        del mine
        res = None
        if rank == 0:
            alone = MultiKrumDefense({"num_malicious": f, "multi_k": mk})
            t1 = time_steps(lambda: alone.aggregate(full, ne), 5, 2, False) / 5
            res = (t1, alone.selected_clients == sel)
        torch.cuda.synchronize()
        dist.barrier()
        t1, same_sel = _bcast(res, dist_on)
        c4["t1_ms_rank0_alone"] = t1 * 1e3
        c4["efficiency"] = t1 / (world * t)
        c4["selection_identical_to_unsharded"] = bool(same_sel)
    else:
        c4["efficiency"] = 1.0
        del mine
    del full
    if with_cpu:
        c4["cpu_baseline"] = cpu_krum_extrapolated(n, d)
    out["c4"] = c4
    # ---------------- C5: 500 x 11,689,512, FLTrust and gradient clipping ----------------
    n, d = 500, 11_689_512
    ne = [50 + (i % 7) for i in range(n)]
    g0_full = torch.randn(d, device="cuda", generator=torch.Generator(device="cuda").manual_seed(5)) * 1e-3

    def make_c5():
        X = synth_rows(n, d, n // 10, 51, theta_sigma=0.0, spread=0.3)
        X[n // 10:] += 0.7 * g0_full
        return X

    def c5_round(ups, group, k, w, on):
        fl = Fixed({"process_group": group})
        tf = time_steps(lambda: fl.aggregate(ups, ne, global_params=[g0_full]), k, w, on) / k
        uu = engine.Updates(ups)
        clip = float(np.median(np.sqrt(engine.row_stats(uu, group=group)["sumsq"])))
        uu.close()
        gc = GradientClippingDefense({"clip_norm": clip, "clip_type": "l2", "process_group": group})
        tc = time_steps(lambda: gc.aggregate(ups, ne), k, w, on) / k
        ok = bool(np.all(np.array(fl.trust_scores[n // 10:]) > 0.1)) and 0 < gc.clipped_count < n
        return tf, tc, ok, (fl.detect_malicious(None, None), gc.clipped_count)

    mine, full, (lo, hi), t_scatter = _ingest(make_c5, n, d, rank, world, dist_on, comm)
    tf, tc, ok, det = c5_round(mine, comm, steps, 3, dist_on)
    alg = 4.0 * (hi - lo) * (2 * n + 2)
    c5 = {"workload": "C5: 500 clients x 11,689,512 params, FLTrust (injected g0) and GradientClipping L2, one round each",
          "n_gpus": world, "cols_per_gpu": hi - lo, "steps": steps, "sane": ok,
          "fltrust": {"api": "FLTrustDefense.aggregate(process_group=flagg.Comm, global_params=g0)", "ms": tf * 1e3,
                      "GBps": 4.0 * n * d / tf / 1e9, "roofline_frac_hbm": alg / tf / 1e9 / hbm},
          "gradient_clipping": {"api": "GradientClippingDefense.aggregate(process_group=flagg.Comm)", "ms": tc * 1e3,
                                "GBps": 4.0 * n * d / tc / 1e9, "roofline_frac_hbm": 4.0 * (hi - lo) * (2 * n + 1) / tc / 1e9 / hbm},
          "algorithmic_bytes_per_gpu": alg}
    if world > 1:
        c5["ingest_scatter"] = {"ms": t_scatter * 1e3, "GBps": 4.0 * n * (d - d // world) / t_scatter / 1e9}
        del mine
        res = None
        if rank == 0:
            t1f, t1c, _, det1 = c5_round(full, None, 5, 2, False)
            res = (t1f, t1c, det1 == det)
        torch.cuda.synchronize()
        dist.barrier()
        t1f, t1c, same_det = _bcast(res, dist_on)
        c5["fltrust"]["t1_ms_rank0_alone"], c5["gradient_clipping"]["t1_ms_rank0_alone"] = t1f * 1e3, t1c * 1e3
        c5["fltrust"]["efficiency"] = t1f / (world * tf)
        c5["gradient_clipping"]["efficiency"] = t1c / (world * tc)
        c5["efficiency"] = c5["fltrust"]["efficiency"]
        c5["detections_identical_to_unsharded"] = bool(same_det)
    else:
        c5["efficiency"] = 1.0
        del mine
    del full
    if with_cpu:
        c5["cpu_baseline"] = cpu_fltrust_sample(n, d, 1_000_000)
    del g0_full
    out["c5"] = c5
    return out


def run_reference(args, rank):
    """The reference's CPU implementation of the headline workload on the host cores, the SAME
    100 x 25M step as the GPU arm (no column sample); --steps / --warmup are honoured unless the
    projected run exceeds REF_BUDGET_S, in which case the counts used are stated in the line."""
    if rank != 0:
        return
    gbs, dt, cores, steps, warm = cpu_reference_median(args.steps, args.warmup, budget_s=REF_BUDGET_S)
    sample = f"full step: {N_CLIENTS} clients x {D_PARAMS} columns (torch CPU, all host threads), {dt:.2f} s/step"
    clamp = None
    if steps != args.steps or warm != args.warmup:
        clamp = f"ran {steps} steps after {warm} warm-ups (requested {args.steps}/{args.warmup}) to stay under {REF_BUDGET_S:.0f} s"
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": dict(CONFIG), "clamp": clamp,
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def run_flagg(args, rank, world, local_rank):
    import torch.distributed as dist
    from flagg import MEDIAN, engine
    from flagg._lib import check, lib
    dist_on = world > 1
    torch.cuda.set_device(local_rank)
    if dist_on:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    # the sharded defenses' one collective goes through the C-ABI communicator (flagg_comm: NCCL), its id over torch.distributed
    comm = engine.Comm.from_torch() if dist_on else None
    parity = sharded_parity_check(comm, rank, world) if dist_on else None

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    n, d = N_CLIENTS, D_PARAMS
    X = synth_rows(n, d, 10, 1234 + rank)
    u = engine.Updates([[X[i]] for i in range(n)])
    out = torch.empty(d, dtype=torch.float32, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    view = u.view()

    def step():
        check(lib().flagg_colwise_select(view, MEDIAN, 0, out.data_ptr(), stream), "flagg_colwise_select")

    step()
    torch.cuda.synchronize()
    sampler.mark()
    t = time_steps(step, args.steps, args.warmup, dist_on)
    sampler.mark()
    clocks = sampler.stop() if rank == 0 else None
    ms = t / args.steps * 1e3
    value = world * 4.0 * n * d * args.steps / t / 1e9
    hbm, which = peaks()
    alg_bytes = 4.0 * d * (n + 1)
    achieved = alg_bytes / (t / args.steps) / 1e9
    launches = args.steps

    plugin = plugin_device_median(X, 10, dist_on, world) if not args.no_extras else None
    e2e = e2e_median(X, out, max(1, min(args.steps, 3)), dist_on, world) if not args.no_e2e else None
    u.close()
    del X, u, out
    torch.cuda.empty_cache()
    with_cpu = world == 1 and not args.no_cpu
    extras, strong, pageable, plugin_r50 = {}, None, None, None
    if not args.no_extras:
        if rank == 0:
            plugin_r50 = plugin_device_resnet50(10)
            pageable = e2e_pageable() if not args.no_e2e else None
            extras = c2_defenses(False, with_cpu)
        if dist_on:
            dist.barrier()
        strong = strong_scaling(rank, world, dist_on, comm, max(10, min(args.steps, 20)), with_cpu)
    if comm is not None:
        comm.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic (Byzantine rows: the reference's ModelReplacementAttack through the attack mirror)",
            "config": dict(CONFIG), "clients_x_params_per_s": world * n * d * args.steps / t,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                         "traffic": ncu_traffic("colselect_bulk_kernel<100,median>"), "peak_source": which,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel": "colselect_bulk_kernel<100, MEDIAN_EXACT, 256> (2-D TMA boxes of 256 columns x 100 rows; the input is one dense matrix)"},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks,
            "plugin_device": plugin, "plugin_device_resnet50": plugin_r50, "e2e_pageable": pageable, "strong": strong, "defenses": extras,
        }
        if dist_on:
            line["sharded_parity"] = parity
        if with_cpu:
            gbs, dt, cores, _, _ = cpu_reference_median(1, 1)
            line["cpu_baseline"] = {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port",
                                    "sample": f"full step {n} clients x {d} columns, 1 step after 1 warm-up "
                                              f"(oracle port of torch.stack+torch.median, {dt:.2f} s/step)"}
        emit(line)
    if dist_on:
        dist.barrier()
        dist.destroy_process_group()


_REAL_STDOUT = None


def quiet_stdout():
    """Libraries (NCCL's version banner) print to fd 1; the contract is ONE JSON line on stdout.
    Everything else is sent to stderr and the line is written to the saved descriptor."""
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _REAL_STDOUT is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_REAL_STDOUT, data)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="flagg", choices=["flagg", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="headline only (used for ncu captures)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    quiet_stdout()
    args.warmup = max(args.warmup, 3) if args.impl == "flagg" else args.warmup
    if args.impl == "reference":
        run_reference(args, rank)
        return
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the flagg arm has no CPU fallback")
    run_flagg(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
