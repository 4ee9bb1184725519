"""Multi-GPU host logic on CPU: column sharding + one all-reduce of the partial Gram /
partial row sums gives every rank the unsharded statistics (world_size 2, gloo).  The
per-shard partials come from the oracle here; on GPUs they come from K2/K4."""
import os
import socket

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import oracle64


def test_shard_bounds_cover_and_align():
    from flagg.sharding import shard_bounds
    for d in [1, 1023, 1024, 25_000_000, 11_689_512, 421_642]:
        for w in [1, 2, 4, 8]:
            b = shard_bounds(d, w)
            assert b[0][0] == 0 and b[-1][1] == d
            for (l0, h0), (l1, h1) in zip(b, b[1:]):
                assert h0 == l1 and l0 <= h0
            assert all(lo % 1024 == 0 or lo == d for lo, _ in b)


def test_shard_update_lists_views():
    from flagg.sharding import shard_update_lists
    ups = [[torch.arange(3000.).reshape(30, 100) + i, torch.arange(500.) - i] for i in range(3)]
    parts = [shard_update_lists(ups, r, 2) for r in range(2)]
    for i in range(3):
        full = torch.cat([t.flatten() for t in ups[i]])
        got = torch.cat([t for r in range(2) for t in parts[r][0][i]])
        assert torch.equal(full, got)
    assert parts[0][0][0][0].data_ptr() == ups[0][0].data_ptr()      # views, no copy


def _worker(rank, world, port, X, f, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from flagg.sharding import shard_update_lists
    from flagg.engine import all_reduce
    ups = [[torch.from_numpy(X[i])] for i in range(X.shape[0])]
    mine, (lo, hi) = shard_update_lists(ups, rank, world)
    Xs = np.stack([torch.cat(r).numpy() if r else np.zeros(0, np.float32) for r in mine])
    Xc = Xs.astype(np.float64) - Xs[0].astype(np.float64)
    G = torch.from_numpy(Xc @ Xc.T)
    stats = torch.from_numpy(oracle64.row_sumsq64(Xs))
    all_reduce(G, None)
    all_reduce(stats, None)
    if rank == 0:
        out.put((G.numpy(), stats.numpy()))
    dist.barrier()
    dist.destroy_process_group()


def test_gram_allreduce_world2_gloo():
    rng = np.random.RandomState(0)
    n, d, f = 12, 5000, 2
    X = (rng.normal(0, 0.05, d)[None, :] + rng.normal(0, 1e-3, (n, d))).astype(np.float32)
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, X, f, q)) for r in range(2)]
    [p.start() for p in procs]
    G, ss = q.get()
    [p.join(60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    g = np.diag(G)
    D = np.sqrt(np.maximum(g[:, None] + g[None, :] - 2 * G, 0))
    np.testing.assert_allclose(D, oracle64.distances64(X), rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(ss, oracle64.row_sumsq64(X), rtol=1e-12)
    assert np.argsort(oracle64.krum_scores64(D, f)).tolist() == np.argsort(oracle64.krum_scores64(oracle64.distances64(X), f)).tolist()
