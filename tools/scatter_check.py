#!/usr/bin/env python
"""Ingest scatter rate under torchrun: rank 0 holds n x d, every rank receives its column shard
(flagg_comm_scatter_columns).   torchrun --nproc-per-node N tools/scatter_check.py [n] [d]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "multimodal-fl-security_b200")]
import torch
import torch.distributed as dist
from flagg import Comm

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 25_000_000
rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
comm = Comm.from_torch()
X = torch.randn((n, d), device="cuda") if rank == 0 else None
ups = [[X[i]] for i in range(n)] if rank == 0 else None
for it in range(3):
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    mine, (lo, hi) = comm.scatter_columns(ups, n, [d], root=0)
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1)], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        sent = 4.0 * n * (d - (hi - lo))
        print(f"N={world} n={n} d={d} scatter {t.item():8.1f} ms = {sent / t.item() / 1e6:7.1f} GB/s off rank 0 "
              f"(NCCL_MIN_P2P_NCHANNELS={os.environ.get('NCCL_MIN_P2P_NCHANNELS')}, NCCL_P2P_NET_CHUNKSIZE={os.environ.get('NCCL_P2P_NET_CHUNKSIZE')})", flush=True)
    del mine
dist.barrier()
comm.close()
dist.destroy_process_group()
