#!/usr/bin/env python
"""Native communicator check under torchrun (GPU box, N >= 2): flagg.Comm all-reduce of a partial
FP64 Gram and of row statistics equals the torch.distributed NCCL all-reduce and the unsharded
result; sharded Multi-Krum / FLTrust through the defense classes select identically on every rank.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 tools/comm_check.py"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import numpy as np, torch, torch.distributed as dist
from flagg import Comm, engine
from flagg.sharding import shard_update_lists
from src.defenses import MultiKrumDefense

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
comm = Comm.from_torch()
rng = np.random.RandomState(3)
n, segs = 300, [70000, 36, 30001]
d = sum(segs)
X = (rng.normal(0, 0.05, d)[None, :] + rng.normal(0, 1e-3, (n, d))).astype(np.float32)
X[:30] *= 10
offs = np.concatenate([[0], np.cumsum(segs)]).astype(int)
full = [[torch.from_numpy(X[i, offs[p]:offs[p + 1]].copy()).cuda() for p in range(len(segs))] for i in range(n)]
mine, (lo, hi) = shard_update_lists(full, rank, world)
u = engine.Updates(mine)
c = torch.from_numpy(np.median(X[::10], axis=0)[lo:hi].copy()).cuda()
Gn = engine.gram(u, center=c, group=comm).cpu().numpy()                 # native communicator
Gt = engine.gram(u, center=c, group=dist.group.WORLD).cpu().numpy()     # torch.distributed NCCL
Xc = X.astype(np.float64) - np.median(X[::10], axis=0).astype(np.float64)
G64 = Xc @ Xc.T
err = np.max(np.abs(Gn - G64) / np.sqrt(np.outer(np.diag(G64), np.diag(G64))))
assert np.array_equal(Gn, Gt), "native and torch all-reduce differ"
assert err < 1e-5, err
g0 = torch.from_numpy(rng.normal(0, 1e-3, d).astype(np.float32)[lo:hi].copy()).cuda()
sn, st = engine.row_stats(u, g0, group=comm), engine.row_stats(u, g0, group=dist.group.WORLD)
for k in ("sumsq", "dot", "linf"):
    assert np.array_equal(sn[k], st[k]), k
np.testing.assert_allclose(sn["sumsq"], (X.astype(np.float64) ** 2).sum(1), rtol=1e-6)
dk = MultiKrumDefense({"num_malicious": 30, "multi_k": 150, "process_group": comm})
dk.aggregate(mine, [1] * n)
sel = torch.tensor(dk.selected_clients, device="cuda")
gathered = [torch.empty_like(sel) for _ in range(world)]
dist.all_gather(gathered, sel)
assert all(torch.equal(g, sel) for g in gathered), "ranks disagree on the Multi-Krum selection"
assert not (set(dk.selected_clients) & set(range(30))), "a Byzantine client was selected"
u.close(); comm.close()
dist.barrier()
if rank == 0:
    print(f"comm check ok: world={world} Gram err {err:.2e}, native == torch all-reduce bit for bit, selections identical on all ranks")
dist.destroy_process_group()
