"""Host side of the engine: turns the reference's ``List[List[Tensor]]`` into a
flagg_view (pointer table, no packing copy) and launches the kernels of libflagg.so on
the caller's current CUDA stream.  torch is used for device memory and streams only.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import FlaggError, MEDIAN, TRIMMED_MEAN, check, lib


def _require_cuda() -> None:
    if not torch.cuda.is_available():
        raise FlaggError("flagg needs a CUDA device (sm_100a); there is no CPU fallback")


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


class Comm:
    """Native communicator of the C-ABI (flagg_comm: NCCL over NVLink), one per process / GPU.
    The only collectives the parameter-sharded path has: the all-reduce of a partial FP64 Gram
    matrix (Krum) and of partial row statistics (FLTrust, clipping, norm bounding).
    ``Comm.from_torch()`` takes rank / world size from torch.distributed and lets rank 0's NCCL
    id travel over that group (gloo or nccl); ``Comm(nranks, rank, id)`` takes them directly."""

    def __init__(self, nranks: int, rank: int, unique_id: bytes):
        _require_cuda()
        if len(unique_id) != 128:
            raise ValueError("the NCCL unique id is 128 bytes")
        self._id = C.create_string_buffer(bytes(unique_id), 128)
        h = C.c_void_p()
        check(lib().flagg_comm_create(C.byref(h), int(nranks), int(rank), self._id), "flagg_comm_create")
        self._h, self.nranks, self.rank = h, int(nranks), int(rank)

    @staticmethod
    def unique_id() -> bytes:
        buf = C.create_string_buffer(128)
        check(lib().flagg_comm_unique_id(buf), "flagg_comm_unique_id")
        return buf.raw

    @classmethod
    def from_torch(cls, group=None) -> "Comm":
        import torch.distributed as dist
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [cls.unique_id() if rank == 0 else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls(world, rank, box[0])

    def all_reduce(self, t: torch.Tensor, op: str = "sum") -> torch.Tensor:
        if not (t.is_cuda and t.is_contiguous()):
            raise ValueError("all_reduce needs a contiguous device tensor")
        code = {"sum": 0, "max": 1}[op]
        if t.dtype == torch.float64:
            check(lib().flagg_allreduce_f64(self._h, t.data_ptr(), t.numel(), code, _stream()), "flagg_allreduce_f64")
        elif t.dtype == torch.float32:
            check(lib().flagg_allreduce_f32(self._h, t.data_ptr(), t.numel(), code, _stream()), "flagg_allreduce_f32")
        else:
            raise ValueError(f"all_reduce: unsupported dtype {t.dtype}")
        return t

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().flagg_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def all_reduce(t: torch.Tensor, group, op: str = "sum") -> torch.Tensor:
    """In place over ``group``: a flagg ``Comm`` (native NCCL communicator) or a torch.distributed
    process group (any backend -- the CPU tests drive the sharded host logic over gloo)."""
    if isinstance(group, Comm):
        return group.all_reduce(t, op)
    import torch.distributed as dist
    dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX, group=group)
    return t


class Updates:
    """n client updates of P tensors each, as the defenses receive them
    (base_defense.py:31-35).  Device tensors are used in place; host tensors stay on the
    host until a kernel needs them (``resident()`` uploads them once)."""

    def __init__(self, client_updates: Sequence[Sequence[torch.Tensor]]):
        if len(client_updates) == 0:
            raise ValueError("client_updates is empty")
        self.n = len(client_updates)
        self.P = len(client_updates[0])
        if self.P == 0:
            raise ValueError("client updates hold no tensors")
        first = client_updates[0]
        self.shapes = [tuple(t.shape) for t in first]
        self.numel = [int(t.numel()) for t in first]
        self.in_dtypes = [t.dtype for t in first]
        self.d = int(sum(self.numel))
        self.on_host = not first[0].is_cuda
        self.device = first[0].device
        rows = []
        numel, on_host, f32 = self.numel, self.on_host, torch.float32   # this loop is the fixed cost of a call at small shapes
        for i, upd in enumerate(client_updates):
            if len(upd) != self.P:
                raise ValueError(f"client {i} has {len(upd)} tensors, expected {self.P}")
            row = []
            push = row.append
            for p, t in enumerate(upd):
                if t.numel() != numel[p]:
                    raise ValueError(f"client {i} tensor {p}: {tuple(t.shape)} != {self.shapes[p]}")
                if t.is_cuda == on_host:
                    raise ValueError("client updates mix host and device tensors")
                if t.requires_grad:
                    t = t.detach()
                # every reference defense .float()s on flatten/stack (krum.py:57, trimmed_mean.py:79)
                if t.dtype is not f32 or not t.is_contiguous():
                    t = t.to(f32).contiguous()
                push(t)
            rows.append(row)
        self.rows = rows          # keeps every tensor alive while a view points at it
        self._view = None
        self._dense = None

    # -- views ---------------------------------------------------------------------
    def _ptr_tables(self):
        seg = (C.c_int64 * self.P)(*self.numel)
        ptrs = (C.c_void_p * (self.n * self.P))()
        k = 0
        for row in self.rows:
            for t in row:
                ptrs[k] = t.data_ptr()
                k += 1
        return seg, ptrs

    def resident(self) -> "Updates":
        """Device-resident version of these updates (uploads host rows into one dense
        [n, ld] matrix with flagg_upload_rows; a no-op for device inputs)."""
        if not self.on_host:
            return self
        if self._dense is None:
            _require_cuda()
            ld = (self.d + 3) & ~3
            X = torch.empty((self.n, ld), dtype=torch.float32, device="cuda")
            seg, ptrs = self._ptr_tables()
            check(lib().flagg_upload_rows(self.n, self.P, seg, ptrs, X.data_ptr(), ld, _stream()), "flagg_upload_rows")
            offs = np.concatenate([[0], np.cumsum(self.numel)])
            lists = [[X[i, offs[p]:offs[p + 1]] for p in range(self.P)] for i in range(self.n)]
            dev = Updates(lists)
            dev.shapes, dev.in_dtypes = self.shapes, self.in_dtypes
            dev._keep = X
            self._dense = dev
        return self._dense

    def view(self) -> int:
        if self.on_host:
            raise FlaggError("internal: view() on host-resident updates")
        if self._view is None:
            _require_cuda()
            seg, ptrs = self._ptr_tables()
            h = C.c_void_p()
            with torch.cuda.device(self.device):
                check(lib().flagg_view_create(C.byref(h), self.n, self.P, seg, ptrs, _stream()), "flagg_view_create")
            self._view = h
        return self._view

    def close(self) -> None:
        if self._view is not None:
            lib().flagg_view_destroy(self._view)
            self._view = None
        if self._dense is not None:
            self._dense.close()
            self._dense = None
        sub = getattr(self, "_center_sub", None)
        if sub is not None:
            self._center_sub = None
            if sub is not self:
                sub.close()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # -- result shaping --------------------------------------------------------------
    def split(self, flat: torch.Tensor, to_host: bool, dtypes=None) -> List[torch.Tensor]:
        """Flat [d] result -> list of P tensors shaped like the inputs, on the inputs' device."""
        if to_host and flat.is_cuda:
            flat = flat.cpu()
        out, off = [], 0
        for p in range(self.P):
            t = flat[off:off + self.numel[p]].reshape(self.shapes[p])
            if dtypes is not None and dtypes[p] != t.dtype:
                t = t.to(dtypes[p])
            out.append(t)
            off += self.numel[p]
        return out


# ---------------------------------------------------------------------------------------
# kernels
# ---------------------------------------------------------------------------------------
def colwise_select(u: Updates, mode: int, k: int = 0) -> torch.Tensor:
    """K1.  Returns the flat fp32 result [d] on the inputs' device (host inputs are
    streamed through the GPU by flagg_colwise_select_host and come back on the host)."""
    _require_cuda()
    if u.on_host:
        out = torch.empty(u.d, dtype=torch.float32)
        seg, ptrs = u._ptr_tables()
        check(lib().flagg_colwise_select_host(u.n, u.P, seg, ptrs, mode, k, out.data_ptr(), 0),
              "flagg_colwise_select_host")
        return out
    with torch.cuda.device(u.device):
        out = torch.empty(u.d, dtype=torch.float32, device=u.device)
        check(lib().flagg_colwise_select(u.view(), mode, k, out.data_ptr(), _stream()), "flagg_colwise_select")
    return out


def row_stats(u: Updates, g0_flat: Optional[torch.Tensor] = None, group=None, centered: bool = False) -> dict:
    """K2.  One pass over X: per-client sum of squares, dot with g0 (FP64) and max |x|
    (of x - g0 when ``centered``).
    With ``group`` (a torch.distributed process group; X sharded by parameters) the
    partial sums are all-reduced so every rank gets whole-vector statistics."""
    u = u.resident()
    with torch.cuda.device(u.device):
        n = u.n
        buf = torch.zeros(2 * n + 1, dtype=torch.float64, device=u.device)
        linf = torch.zeros(n, dtype=torch.float32, device=u.device)
        g_ptr = None
        if g0_flat is not None:
            g0_flat = g0_flat.to(device=u.device, dtype=torch.float32).contiguous()
            if g0_flat.numel() != u.d:
                raise ValueError("g0 has the wrong number of elements")
            g_ptr = g0_flat.data_ptr()
        p = buf.data_ptr()
        check(lib().flagg_row_reduce(u.view(), g_ptr, 1 if centered else 0, p, (p + 8 * n) if g_ptr else None, linf.data_ptr(),
                                     (p + 16 * n) if g_ptr else None, _stream()), "flagg_row_reduce")
        if group is not None:
            all_reduce(buf, group, "sum")
            all_reduce(linf, group, "max")
        h = buf.cpu().numpy()
        return dict(sumsq=h[:n].copy(), dot=h[n:2 * n].copy(), g0_sumsq=float(h[2 * n]),
                    linf=linf.cpu().numpy())


def row_combine(u: Updates, rows, w1, w2, divisor: float) -> torch.Tensor:
    """K3.  Flat fp32 [d] = (sum_j w2_j*(w1_j*x_rows[j])) / divisor, reference order."""
    u = u.resident()
    m = len(rows) if rows is not None else u.n
    def arr(x, ct):
        if x is None:
            return None
        a = np.ascontiguousarray(x, dtype=np.float32 if ct is C.c_float else np.int32)
        if a.size != m:
            raise ValueError("weight / row arrays must have one entry per combined row")
        return a.ctypes.data_as(C.POINTER(ct)), a
    r = arr(rows, C.c_int32)
    a1 = arr(w1, C.c_float)
    a2 = arr(w2, C.c_float)
    with torch.cuda.device(u.device):
        out = torch.empty(u.d, dtype=torch.float32, device=u.device)
        check(lib().flagg_row_combine(u.view(), m, r[0] if r else None, a1[0] if a1 else None,
                                      a2[0] if a2 else None, float(np.float32(divisor)), out.data_ptr(), _stream()),
              "flagg_row_combine")
    return out


CENTER_ROWS = 31   # rows whose coordinate-wise median centres the Gram matrix


def robust_center(u: Updates) -> torch.Tensor:
    """A vector inside the benign cluster: the coordinate-wise median (K1) of up to 31
    evenly spaced clients.  It only conditions the Gram arithmetic (distances are
    translation invariant), so it need not be exact or agreed between ranks: with the
    parameter dimension sharded every rank centres its own columns."""
    u = u.resident()
    sub = getattr(u, "_center_sub", None)
    if sub is None:      # the sub-view (pointer tables, tensor maps) is built once per Updates, like the main view
        idx = sorted(set(int(round(x)) for x in np.linspace(0, u.n - 1, min(u.n, CENTER_ROWS))))
        # up to 31 clients every row is used: the Updates is its own sub-view (no second pointer table / view)
        sub = u._center_sub = u if len(idx) == u.n else Updates([u.rows[i] for i in idx])
    return colwise_select(sub, MEDIAN)


def gram(u: Updates, center: Optional[torch.Tensor] = None, group=None, auto_center: bool = True) -> torch.Tensor:
    """K4.  FP64 centred Gram matrix [n, n] on the device; all-reduced over ``group``
    when the parameter dimension is sharded (partial Grams add).  ``center``: flat fp32
    [d] on the device; by default ``robust_center(u)``."""
    u = u.resident()
    with torch.cuda.device(u.device):
        if center is None and auto_center:
            center = robust_center(u)
        if center is not None:
            center = center.to(device=u.device, dtype=torch.float32).contiguous()
            if center.numel() != u.d:
                raise ValueError("center has the wrong number of elements")
        G = torch.empty((u.n, u.n), dtype=torch.float64, device=u.device)
        check(lib().flagg_gram(u.view(), center.data_ptr() if center is not None else None, 0, G.data_ptr(),
                               _stream()), "flagg_gram")
        if group is not None:
            all_reduce(G, group, "sum")
    return G


def krum_scores(G: torch.Tensor, f: int, want_D: bool = True):
    """K5.  (D [n,n] float64 ndarray or None, scores [n] float64 ndarray)."""
    n = G.shape[0]
    with torch.cuda.device(G.device):
        scores = torch.empty(n, dtype=torch.float64, device=G.device)
        D = torch.empty((n, n), dtype=torch.float64, device=G.device) if want_D else None
        check(lib().flagg_krum_scores(G.data_ptr(), n, f, D.data_ptr() if want_D else None, scores.data_ptr(),
                                      _stream()), "flagg_krum_scores")
        return (D.cpu().numpy() if want_D else None), scores.cpu().numpy()


def device_info() -> dict:
    _require_cuda()
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    check(lib().flagg_device_info(C.byref(a), C.byref(b), C.byref(c)), "flagg_device_info")
    return dict(sm_count=a.value, cc=(b.value, c.value))
