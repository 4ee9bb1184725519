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
CPU_SAMPLE_COLS = 2_000_000


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        with open(p) as fh:
            return float(json.load(fh)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def tf32_peak_tflops():
    """Dense TF32 GEMM rate of this GPU (library GEMM, used only as the roofline denominator)."""
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
    torch.backends.cuda.matmul.allow_tf32 = old
    del a, b
    return 2.0 * 8192 ** 3 / best / 1e9


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
        self.samples, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.samples.append([x.strip() for x in line.split(",")])

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        rows = [r for r in self.samples if len(r) == 6 and r[0].isdigit()]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({nm for r in rows for nm, v in zip(names, r[2:]) if v.lower().startswith("active")})
        return {"sm_mhz": float(np.median([int(r[0]) for r in rows])), "sm_max_mhz": int(rows[0][1]),
                "reasons": reasons, "samples": len(rows)}


def synth_rows(n, d, n_byz, seed, device="cuda"):
    """Rows as run_experiments.py:238 produces them: full weights theta + delta_i; the first
    n_byz rows are model-replacement style outliers theta + 10*delta (model_poisoning.py:65-80)."""
    g = torch.Generator(device=device).manual_seed(seed)
    theta = torch.randn(d, device=device, generator=g) * 0.05
    X = torch.empty((n, d), dtype=torch.float32, device=device)
    for i in range(n):
        X[i] = torch.randn(d, device=device, generator=g).mul_(1e-2 if i < n_byz else 1e-3).add_(theta)
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


def cpu_reference_median(steps, warmup, n=N_CLIENTS, cols=CPU_SAMPLE_COLS):
    """The reference's CPU path for this workload (torch.stack + torch.median(dim=0),
    trimmed_mean.py:158-164) through the oracle port, on a bounded column sample."""
    from oracle import ref_port
    torch.set_num_threads(os.cpu_count() or 1)
    g = torch.Generator().manual_seed(1234)
    theta = torch.randn(cols, generator=g) * 0.05
    rows = [[theta + torch.randn(cols, generator=g) * (1e-2 if i < 10 else 1e-3)] for i in range(n)]
    def step():
        stacked = torch.stack([u[0].float() for u in rows])      # the reference's packing copy is part of its path
        return ref_port.median(stacked)
    for _ in range(warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(steps):
        step()
    dt = (time.perf_counter() - t0) / steps
    return 4.0 * n * cols / dt / 1e9, dt, torch.get_num_threads()


def run_reference(args, rank):
    if rank != 0:
        return
    steps, warm = max(1, min(args.steps, 5)), max(1, min(args.warmup, 2))
    gbs, dt, cores = cpu_reference_median(steps, warm)
    sample = f"{N_CLIENTS} clients x {CPU_SAMPLE_COLS} of {D_PARAMS} columns per step (torch CPU, all host threads)"
    line = {
        "impl": "reference", "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": warm, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C3: 100 clients x 25,000,000 params, coordinate-wise median (bounded column sample)",
                   "n_clients": N_CLIENTS, "n_params": D_PARAMS},
        "cpu_baseline": {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": gbs, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit(line)


def extra_defenses(rank, world, dist_on, group):
    """Other defenses of the path, short runs (not the headline): trimmed mean (C2),
    FLTrust + clipping stages (C5) and the Krum Gram (C2 cohort, C4 shard).  GB/s = 4*n*d/t."""
    from flagg import TRIMMED_MEAN, engine
    hbm, _ = peaks()
    out = {}

    def bench(name, n, d, fn_factory, alg_bytes, steps=3):
        torch.cuda.empty_cache()
        X = synth_rows(n, d, max(1, n // 10), 99 + rank)
        u = engine.Updates([[X[i]] for i in range(n)])
        fn = fn_factory(u, X)
        t = time_steps(fn, steps, 3, dist_on) / steps
        out[name] = {"n": n, "d_per_gpu": d, "ms": t * 1e3, "GBps": world * 4.0 * n * d / t / 1e9,
                     "clients_x_params_per_s": world * n * d / t,
                     "roofline_frac_hbm": (alg_bytes / t / 1e9) / hbm}
        u.close()
        del X, u

    bench("trimmed_mean_C2_20x11.69M_k2", 20, 11_689_512,
          lambda u, X: (lambda: engine.colwise_select(u, TRIMMED_MEAN, 2)), 4.0 * 11_689_512 * 21)
    n5, d5 = 500, 11_689_512
    g0 = torch.randn(d5, device="cuda") * 1e-3

    def fltrust_factory(u, X):
        w = np.ones(n5, dtype=np.float32)
        def step():
            st = engine.row_stats(u, g0, group=group)            # K2 (+ all-reduce of 2n+1 doubles when sharded)
            sc = (np.sqrt(st["g0_sumsq"]) / np.sqrt(st["sumsq"])).astype(np.float32)
            engine.row_combine(u, None, sc, w, float(n5))         # K3
        return step
    bench("fltrust_C5_500x11.69M", n5, d5, fltrust_factory, 4.0 * d5 * (2 * n5 + 2))
    def krum_c2_factory(u, X):
        def step():
            G = engine.gram(u, group=group)         # K4, small-cohort kernel (gram_small.cu)
            engine.krum_scores(G, 5, want_D=False)                # K5
        return step
    bench("krum_gram_C2_20x11.69M_f5", 20, 11_689_512, krum_c2_factory, 4.0 * 20 * 11_689_512, steps=5)
    n4, d4 = 1000, 25_000_000 // 8

    def krum_factory(u, X):
        def step():
            G = engine.gram(u, group=group)         # K4 (+ all-reduce of the n x n FP64 Gram)
            engine.krum_scores(G, 100, want_D=False)              # K5
        return step
    bench("multikrum_gram_C4shard_1000x3.125M_f100", n4, d4, krum_factory, 4.0 * n4 * d4, steps=3)
    gk = out["multikrum_gram_C4shard_1000x3.125M_f100"]
    # tensor roofline of the Gram kernel: algorithmic flops n(n+1)d (unique pairs, mul+add) against
    # exact-FP32-by-3xTF32 peak = measured dense TF32 GEMM rate / 3 (cuBLAS 8192^3, best of 10, same process)
    tf32 = tf32_peak_tflops()
    gk["TFLOPs_alg"] = world * n4 * (n4 + 1) * d4 / (gk["ms"] / 1e3) / 1e12
    gk["roofline"] = {"bound": "tensor", "achieved": gk["TFLOPs_alg"] / world, "peak": tf32 / 3.0, "unit": "TFLOP/s",
                      "frac": gk["TFLOPs_alg"] / world / (tf32 / 3.0), "tf32_dense_measured": tf32,
                      "kernel": "gram_ts2_kernel (tcgen05 cta_group::2 kind::tf32, 256x256 CTA-pair tiles, A from TMEM, 3 products per k-step)"}
    return out


def run_flagg(args, rank, world, local_rank):
    import torch.distributed as dist
    from flagg import MEDIAN, engine
    from src.defenses import MedianDefense
    dist_on = world > 1
    torch.cuda.set_device(local_rank)
    if dist_on:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    n, d = N_CLIENTS, D_PARAMS
    X = synth_rows(n, d, 10, 1234 + rank)
    u = engine.Updates([[X[i]] for i in range(n)])
    out = torch.empty(d, dtype=torch.float32, device="cuda")
    from flagg._lib import check, lib
    stream = torch.cuda.current_stream().cuda_stream
    view = u.view()

    def step():
        check(lib().flagg_colwise_select(view, MEDIAN, 0, out.data_ptr(), stream), "flagg_colwise_select")

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    t = time_steps(step, args.steps, args.warmup, dist_on)
    clocks = sampler.stop() if rank == 0 else None
    ms = t / args.steps * 1e3
    value = world * 4.0 * n * d * args.steps / t / 1e9
    hbm, which = peaks()
    alg_bytes = 4.0 * d * (n + 1)
    achieved = alg_bytes / (t / args.steps) / 1e9

    # end to end through the plugin API with pinned HOST tensors in and host tensors out
    # (Flower path, robust_server.py:99-125): H2D of all n rows and D2H of the result inside the timed region
    e2e, same, e2e_steps, pcie = None, None, 0, None
    if not args.no_e2e:
        host = torch.empty((n, d), dtype=torch.float32, pin_memory=True)
        host.copy_(X)
        torch.cuda.synchronize()
        # plain pinned H2D rate of this box = the ceiling of any host-input path
        t0 = time.perf_counter()
        X.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        pcie = 4.0 * n * d / (time.perf_counter() - t0) / 1e9
        host_lists = [[host[i]] for i in range(n)]
        defense = MedianDefense({})
        e2e_steps = max(1, min(args.steps, 3))
        defense.aggregate(host_lists, [1] * n)
        if dist_on:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            res = defense.aggregate(host_lists, [1] * n)
        torch.cuda.synchronize()
        te = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device="cuda")
        if dist_on:
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
        e2e = world * 4.0 * n * d * e2e_steps / float(te.item()) / 1e9
        same = bool(torch.equal(res[0].view(torch.int32), out.cpu().view(torch.int32)))
        del host, host_lists, res
    u.close()
    del X, u
    # the sharded defenses' one collective goes through the C-ABI communicator (flagg_comm: NCCL), its id over torch.distributed
    comm = engine.Comm.from_torch() if (dist_on and not args.no_extras) else None
    extras = extra_defenses(rank, world, dist_on, comm) if not args.no_extras else {}
    if comm is not None:
        comm.close()

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C3: 100 clients x 25,000,000 params per GPU, coordinate-wise median",
                       "n_clients": n, "n_params_per_gpu": d, "sharding": f"parameter dim x{world}, no collective",
                       "l2_policy": "inputs (10 GB) larger than L2 (126 MB); no flush needed",
                       "clients_x_params_per_s": world * n * d * args.steps / t},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": hbm, "unit": "GB/s", "frac": achieved / hbm,
                         "traffic": ncu_traffic("colselect_bulk_kernel<100,median>"), "peak_source": which,
                         "algorithmic_bytes_per_launch": alg_bytes, "kernel": "colselect_bulk_kernel<100, MEDIAN_EXACT> (2-D TMA feed; the input is one dense matrix)"},
            "e2e": {"value": e2e, "unit": "GB/s", "h2d_bytes_per_step": 4 * n * d, "d2h_bytes_per_step": 4 * d,
                    "steps": e2e_steps, "api": "src.defenses.MedianDefense.aggregate(pinned host tensors)",
                    "matches_device_result": same, "pinned_h2d_memcpy_GBps": pcie},
            "gpu_launches": args.steps, "clocks": clocks, "defenses": extras,
        }
        if world == 1 and not args.no_cpu:
            gbs, dt, cores = cpu_reference_median(1, 1)
            line["cpu_baseline"] = {"value": gbs, "unit": "GB/s", "cores": cores, "kind": "port",
                                    "sample": f"{n} clients x {CPU_SAMPLE_COLS} of {d} columns, 1 step after 1 warm-up "
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
