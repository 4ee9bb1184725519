"""GPU parity, part 5: the parameter-sharded path on real GPUs (one process per GPU, the native
flagg.Comm communicator = NCCL through the C-ABI).  Every rank holds the whole cohort on its own
device, aggregates (a) unsharded and (b) its column shard with ``process_group=comm``; the sharded
call must select exactly what the unsharded one selects and produce that shard of its output.
Skipped on a box with fewer than two GPUs (run it with ``gpurun --gpus 2``)."""
import os
import socket
import traceback

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _rows(n, seg, byz, seed):
    rng = np.random.RandomState(seed)
    d = sum(seg)
    theta = rng.normal(0, 0.05, d)
    sig = 1e-3 * (1 + 0.3 * rng.uniform(-1, 1, n))
    X = theta[None, :] + rng.normal(0, 1, (n, d)) * sig[:, None]
    for b in range(byz):
        X[b] = theta + 10 * (X[b] - theta)
    return X.astype(np.float32)


def _lists(X, seg):
    offs = np.concatenate([[0], np.cumsum(seg)]).astype(int)
    return [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(seg))] for i in range(X.shape[0])]


def _flat(res):
    return torch.cat([t.detach().reshape(-1).float() for t in res])


def _worker(rank, world, port, errq):
    try:
        import sys
        root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
        for p in (root, os.path.join(root, "multimodal-fl-security_b200")):
            if p not in sys.path:
                sys.path.insert(0, p)
        import torch.distributed as dist
        os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        torch.cuda.set_device(rank)
        dist.init_process_group("gloo", rank=rank, world_size=world)       # carries the NCCL id only
        from flagg import Comm
        from flagg.sharding import shard_bounds, shard_update_lists
        from src.defenses import (FLTrustDefense, GeometricMedianDefense, GradientClippingDefense, KrumDefense,
                                  MedianDefense, MultiKrumDefense, NormBoundingDefense, TrimmedMeanDefense)
        comm = Comm.from_torch()

        class Fixed(FLTrustDefense):
            def _compute_server_gradient(self, global_params):
                self.server_gradient = global_params
                return global_params

        def both(cls, cfg, ups, ne, **kw):
            """(unsharded defense, its flat output, sharded defense, the shard's flat output, (lo, hi))"""
            a = cls(dict(cfg))
            full = _flat(a.aggregate(ups, ne, **kw))
            mine, (lo, hi) = shard_update_lists(ups, rank, world)
            b = cls(dict(cfg, process_group=comm))
            part = _flat(b.aggregate(mine, ne, **kw))
            return a, full, b, part, (lo, hi)

        def same_bits(x, y):
            return torch.equal(x.contiguous().view(torch.int32), y.contiguous().view(torch.int32))

        # ---- Krum / Multi-Krum at three cohort sizes (small-cohort, single-CTA and pair Gram kernels) ----
        for n, seg, f, mk in ((12, [3000, 77, 5000], 2, 1), (20, [40_000, 333, 60_000], 4, 10), (100, [50_000, 7, 20_000], 10, 50),
                              (300, [9000, 4100], 30, 150)):
            X = _rows(n, seg, f, 100 + n)
            ups = _lists(X, seg)
            cls = KrumDefense if mk <= 10 else MultiKrumDefense
            a, full, b, part, (lo, hi) = both(cls, {"num_malicious": f, "multi_k": mk}, ups, [1] * n)
            assert a.selected_clients == b.selected_clients, (n, a.selected_clients, b.selected_clients)
            assert a.rejected_clients == b.rejected_clients
            np.testing.assert_allclose(np.array(a.client_scores), np.array(b.client_scores), rtol=1e-6)
            assert same_bits(part, full[lo:hi]), f"Krum n={n}: sharded mean differs from the unsharded shard"
            sel = [None] * world
            dist.all_gather_object(sel, b.selected_clients)
            assert all(s == sel[0] for s in sel), "ranks disagree on the selection"
        # ---- a NaN client under sharding: every rank takes the exact path and rejects it ----
        n, seg, f = 40, [6000, 2000], 5
        X = _rows(n, seg, 0, 7)
        X[3, 100] = np.nan                                # the NaN sits in rank 0's shard only; the all-reduced statistics route every rank
        ups = _lists(X, seg)
        a, full, b, part, (lo, hi) = both(KrumDefense, {"num_malicious": f, "multi_k": 10}, ups, [1] * n)
        assert a.selected_clients == b.selected_clients and b.rejected_clients[-1] == 3
        # ---- FLTrust, clipping, norm bounding: all-reduced norms / dots ----
        n, seg = 50, [30_000, 1234, 20_000]
        X = _rows(n, seg, 5, 11)
        g0 = (np.random.RandomState(12).normal(0, 1e-3, sum(seg))).astype(np.float32)
        X[10:] += 0.7 * g0
        X[5:10] -= g0
        ups = _lists(X, seg)
        g0l = _lists(g0[None, :], seg)[0]
        a, full, b, part, (lo, hi) = both(Fixed, {}, ups, [1] * n, global_params=g0l)
        np.testing.assert_allclose(np.array(a.trust_scores), np.array(b.trust_scores), rtol=0, atol=1e-7)
        assert a.detect_malicious(None, None) == b.detect_malicious(None, None)
        assert torch.allclose(part, full[lo:hi], rtol=1e-6, atol=1e-9)
        norms = np.linalg.norm(X.astype(np.float64), axis=1)
        ne = [10 + i for i in range(n)]
        for ctype, clip in (("l2", float(np.median(norms))), ("linf", float(np.median(np.abs(X).max(axis=1))))):
            a, full, b, part, (lo, hi) = both(GradientClippingDefense, {"clip_norm": clip, "clip_type": ctype}, ups, ne)
            assert a.clipped_count == b.clipped_count
            np.testing.assert_allclose(a.original_norms, b.original_norms, rtol=1e-6)
            assert torch.allclose(part, full[lo:hi], rtol=1e-6, atol=1e-9)
        srt = np.sort(norms)
        a, full, b, part, (lo, hi) = both(NormBoundingDefense, {"max_norm": float((srt[-6] + srt[-5]) / 2), "min_norm": 0.0}, ups, ne)
        assert a.rejected_clients == b.rejected_clients == list(range(5))
        assert same_bits(part, full[lo:hi])
        # ---- geometric median: every rank stops at the same iteration ----
        n, seg = 15, [5000, 3000]
        ups = _lists(_rows(n, seg, 3, 13), seg)
        a, full, b, part, (lo, hi) = both(GeometricMedianDefense, {"max_iters": 60, "tolerance": 1e-6}, ups, [1] * n)
        its = [None] * world
        dist.all_gather_object(its, b.num_iters)
        assert all(i == its[0] for i in its) and abs(a.num_iters - b.num_iters) <= 1
        assert torch.allclose(part, full[lo:hi], rtol=1e-5, atol=1e-9)
        # ---- median / trimmed mean: no collective, the shard is the slice ----
        for cls, cfg in ((MedianDefense, {}), (TrimmedMeanDefense, {"trim_ratio": 0.2})):
            a, full, b, part, (lo, hi) = both(cls, cfg, ups, [1] * n)
            assert same_bits(part, full[lo:hi])
        # ---- an EMPTY shard: d < 1024 * world leaves the last rank without columns; it still joins every collective ----
        n, seg, f = 12, [500, 100], 2
        X = _rows(n, seg, f, 31)
        g0 = (np.random.RandomState(32).normal(0, 1e-3, sum(seg))).astype(np.float32)
        X[f:] += 0.7 * g0
        ups, g0l = _lists(X, seg), _lists(g0[None, :], seg)[0]
        assert shard_bounds(sum(seg), world)[-1][0] == shard_bounds(sum(seg), world)[-1][1]
        a, full, b, part, (lo, hi) = both(KrumDefense, {"num_malicious": f, "multi_k": 4}, ups, [1] * n)
        assert a.selected_clients == b.selected_clients and part.numel() == hi - lo and same_bits(part, full[lo:hi])
        a, full, b, part, (lo, hi) = both(Fixed, {}, ups, [1] * n, global_params=g0l)
        np.testing.assert_allclose(np.array(a.trust_scores), np.array(b.trust_scores), rtol=0, atol=1e-7)
        assert torch.allclose(part, full[lo:hi], rtol=1e-6, atol=1e-9)
        a, full, b, part, (lo, hi) = both(GradientClippingDefense, {"clip_norm": float(np.median(np.linalg.norm(X, axis=1)))}, ups, [3] * n)
        assert a.clipped_count == b.clipped_count and torch.allclose(part, full[lo:hi], rtol=1e-6, atol=1e-9)
        a, full, b, part, (lo, hi) = both(MedianDefense, {}, ups, [1] * n)
        assert same_bits(part, full[lo:hi])
        # ---- ingest: rank 0 holds the whole cohort, every rank receives its column shard over NCCL ----
        n, seg, f = 40, [50_000, 1234, 30_000, 7], 5
        X = _rows(n, seg, f, 21)
        ups = _lists(X, seg) if rank == 0 else None
        for copy_root in (False, True):
            mine, (lo, hi) = comm.scatter_columns(ups, n, seg, root=0, copy_root=copy_root)
            got = torch.stack([torch.cat([t.reshape(-1) for t in row]) for row in mine]).cpu().numpy()
            assert np.array_equal(got.view(np.uint32), X[:, lo:hi].view(np.uint32)), "scattered shard differs from the root's columns"
            b = MultiKrumDefense({"num_malicious": f, "multi_k": 20, "process_group": comm})
            part = _flat(b.aggregate(mine, [1] * n))
            if rank == 0:
                a = MultiKrumDefense({"num_malicious": f, "multi_k": 20})
                full = _flat(a.aggregate(ups, [1] * n))
                assert a.selected_clients == b.selected_clients and same_bits(part, full[lo:hi])
        torch.cuda.synchronize()
        dist.barrier()
        comm.close()
        dist.destroy_process_group()
    except Exception:
        errq.put(f"rank {rank}:\n{traceback.format_exc()}")
        raise


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_sharded_defenses_equal_unsharded_two_gpus():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    errq = ctx.SimpleQueue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, errq)) for r in range(2)]
    [p.start() for p in procs]
    [p.join(150) for p in procs]          # a rank that raised leaves its peer in a collective: do not wait for long
    for p in procs:
        if p.is_alive():
            p.kill()
    errs = []
    while not errq.empty():
        errs.append(errq.get())
    assert not errs and all(p.exitcode == 0 for p in procs), "\n".join(errs) or [p.exitcode for p in procs]
