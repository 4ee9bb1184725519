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
from ._lib import FlaggError, MEDIAN, TRIMMED_MEAN, T_AXPY, T_REPLACE, T_SCALE, T_SIGN, check, lib  # noqa: F401


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

    def scatter_columns(self, client_updates, n: int, numel: Sequence[int], root: int = 0, copy_root: bool = False):
        """Ingest for the sharded path: rank ``root`` holds the whole client updates on its device
        (``client_updates``: List[List[Tensor]]; ignored elsewhere), every rank gets its column shard
        (flagg.sharding.shard_bounds) of every client.  ``n`` clients, ``numel``: elements of each of the P
        tensors (known to every rank -- the model's shapes).  Returns (updates, (lo, hi)) where ``updates``
        is a list of n one-tensor lists over a dense [n, hi - lo] device matrix (on the root: zero-copy
        views of its own rows unless ``copy_root``).  flagg_comm_scatter_columns: ncclSend / ncclRecv per
        (row, peer) over NVLink, enqueued on the current stream."""
        from .sharding import shard_bounds, shard_update_lists
        numel = [int(x) for x in numel]
        P, d = len(numel), int(sum(numel))
        bounds = shard_bounds(d, self.nranks)
        lo, hi = bounds[self.rank]
        los = (C.c_int64 * self.nranks)(*[b[0] for b in bounds])
        his = (C.c_int64 * self.nranks)(*[b[1] for b in bounds])
        seg = (C.c_int64 * P)(*numel)
        ptrs, keep, dst = None, None, None
        if self.rank == root:
            u = Updates(client_updates)
            if u.on_host or u.n != n or u.numel != numel:
                raise ValueError("scatter_columns: the root's updates must be device tensors of the announced shapes")
            _, ptrs = u._ptr_tables()
            keep = u
        if self.rank != root or copy_root:
            dst = torch.empty((n, max(hi - lo, 1)), dtype=torch.float32, device="cuda")
        check(lib().flagg_comm_scatter_columns(self._h, int(root), int(n), P, seg, ptrs, los, his,
                                               dst.data_ptr() if dst is not None else None, max(hi - lo, 1), _stream()),
              "flagg_comm_scatter_columns")
        if dst is not None:
            return [[dst[i, :hi - lo]] for i in range(n)], (lo, hi)
        mine, _ = shard_update_lists(keep.rows, self.rank, self.nranks)
        return mine, (lo, hi)

    def close(self) -> None:
        if getattr(self, "_h", None):
            lib().flagg_comm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def group_rank_world(group):
    """(rank, world) of a flagg ``Comm`` or a torch.distributed group; (0, 1) for None."""
    if group is None:
        return 0, 1
    if isinstance(group, Comm):
        return group.rank, group.nranks
    import torch.distributed as dist
    return dist.get_rank(group), dist.get_world_size(group)


def all_reduce(t: torch.Tensor, group, op: str = "sum") -> torch.Tensor:
    """In place over ``group``: a flagg ``Comm`` (native NCCL communicator) or a torch.distributed
    process group (any backend -- the CPU tests drive the sharded host logic over gloo)."""
    if isinstance(group, Comm):
        return group.all_reduce(t, op)
    import torch.distributed as dist
    dist.all_reduce(t, op=dist.ReduceOp.SUM if op == "sum" else dist.ReduceOp.MAX, group=group)
    return t


_HOSTWALK = False     # flagg/_hostwalk*.so (csrc/hostwalk.cpp): resolved on first use; None when it was not built


def _hostwalk():
    global _HOSTWALK
    if _HOSTWALK is False:
        try:
            from . import _hostwalk as mod
            _HOSTWALK = mod
        except ImportError:
            _HOSTWALK = None
    return _HOSTWALK


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
        self._view = None
        self._dense = None
        self._ptrs = None
        hw = _hostwalk()
        if hw is not None and isinstance(client_updates, (list, tuple)):
            # native walk: validates the n x P handles (fp32, contiguous, grad-free, one device, consistent sizes) and
            # fills the pointer table in one pass, ~20 ns per tensor instead of ~0.8 us; anything else -- a conversion
            # to make, an error to report -- takes the general loop below, which words the error
            table = (C.c_void_p * (self.n * self.P))()
            status, n, P, on_host, _dev = hw.walk(client_updates, C.addressof(table))
            if status == 0:
                self.shapes = [tuple(t.shape) for t in first]
                self.numel = [int(t.numel()) for t in first]
                self.in_dtypes = [torch.float32] * self.P
                self.d = int(sum(self.numel))
                self.on_host = bool(on_host)
                self.device = first[0].device
                self.rows = client_updates          # the caller's own lists: they keep every tensor alive
                self._ptrs = table
                return
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

    # -- views ---------------------------------------------------------------------
    def _ptr_tables(self):
        seg = (C.c_int64 * self.P)(*self.numel)
        if self._ptrs is not None:
            return seg, self._ptrs
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

    # -- round epilogue: results written straight into the caller's parameter tensors ------------
    def check_outputs(self, outputs) -> list:
        """Validate ``outputs`` (e.g. ``list(global_model.parameters())``) as the destination of a
        [d] result of these updates: P contiguous fp32 device tensors of the inputs' sizes."""
        outs = [o.data if isinstance(o, torch.nn.Parameter) else o for o in outputs]
        if self.on_host:
            raise ValueError("'output_params' needs device-resident client updates")
        if len(outs) != self.P:
            raise ValueError(f"'output_params' holds {len(outs)} tensors, the updates {self.P}")
        for p, o in enumerate(outs):
            if not (o.is_cuda and o.device == self.device and o.dtype == torch.float32 and o.is_contiguous()
                    and o.numel() == self.numel[p]):
                raise ValueError(f"'output_params'[{p}] must be a contiguous fp32 tensor of {self.numel[p]} elements on {self.device}")
        return outs

    def _bind(self, outs) -> None:
        arr = (C.c_void_p * self.P)(*[o.data_ptr() for o in outs]) if outs is not None else None
        check(lib().flagg_view_set_outputs(self.view(), arr, _stream()), "flagg_view_set_outputs")

    def scatter(self, flat: torch.Tensor, outs) -> None:
        """flat [d] -> the P output tensors (one launch; used where the producing kernel could not
        write them itself)."""
        off, xs = 0, []
        for p in range(self.P):
            xs.append(flat[off:off + self.numel[p]])
            off += self.numel[p]
        segments_transform(T_SCALE, 1.0, xs, None, outs)

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
            if sub is not True:
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
        # one split call (P views from C++), then one view() per tensor: at P = 161 the slice + reshape loop this
        # replaces cost more than the native walk over all 16,100 input handles
        parts = flat.split(self.numel) if self.P > 1 else (flat,)
        out = [t.view(s) for t, s in zip(parts, self.shapes)]
        if dtypes is not None:
            out = [t if d == t.dtype else t.to(d) for t, d in zip(out, dtypes)]
        return out


# ---------------------------------------------------------------------------------------
# kernels
# ---------------------------------------------------------------------------------------
def colwise_select(u: Updates, mode: int, k: int = 0, outputs=None) -> Optional[torch.Tensor]:
    """K1.  Returns the flat fp32 result [d] on the inputs' device (host inputs are
    streamed through the GPU by flagg_colwise_select_host and come back on the host).
    With ``outputs`` (P device tensors, see Updates.check_outputs) the kernel writes them directly
    and None is returned."""
    _require_cuda()
    if outputs is not None:
        outs = u.check_outputs(outputs)
        with torch.cuda.device(u.device):
            u._bind(outs)
            try:
                check(lib().flagg_colwise_select(u.view(), mode, k, None, _stream()), "flagg_colwise_select")
            finally:
                u._bind(None)
        return None
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
            g_ptr = g0_flat.data_ptr() or None          # empty column shard: zeros are this rank's partial sums
        p = buf.data_ptr()
        check(lib().flagg_row_reduce(u.view(), g_ptr, 1 if centered else 0, p, (p + 8 * n) if g_ptr else None, linf.data_ptr(),
                                     (p + 16 * n) if g_ptr else None, _stream()), "flagg_row_reduce")
        if group is not None:
            all_reduce(buf, group, "sum")
            all_reduce(linf, group, "max")
        h = buf.cpu().numpy()
        return dict(sumsq=h[:n].copy(), dot=h[n:2 * n].copy(), g0_sumsq=float(h[2 * n]),
                    linf=linf.cpu().numpy())


def row_combine(u: Updates, rows, w1, w2, divisor: float, outputs=None) -> Optional[torch.Tensor]:
    """K3.  Flat fp32 [d] = (sum_j w2_j*(w1_j*x_rows[j])) / divisor, reference order.
    With ``outputs`` the kernel writes the P tensors directly and None is returned."""
    outs = u.check_outputs(outputs) if outputs is not None else None
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
        out = torch.empty(u.d, dtype=torch.float32, device=u.device) if outs is None else None
        if outs is not None:
            u._bind(outs)
        try:
            check(lib().flagg_row_combine(u.view(), m, r[0] if r else None, a1[0] if a1 else None,
                                          a2[0] if a2 else None, float(np.float32(divisor)),
                                          out.data_ptr() if out is not None else None, _stream()),
                  "flagg_row_combine")
        finally:
            if outs is not None:
                u._bind(None)
    return out


W_FLTRUST, W_CLIP, W_DPSGD, W_NORMBOUND = 0, 1, 2, 3


def weighted_round(u: Updates, rule: int, num_examples=None, g0_flat: Optional[torch.Tensor] = None, p0: float = 0.0,
                   p1: float = 0.0, use_linf: bool = False, group=None, outputs=None):
    """One clipping / trust round with the host out of the middle: K2 (row statistics) -> all-reduce when
    sharded -> flagg_round_weights (the reference's per-client arithmetic on the device) -> K3 with
    device-resident rows / weights / divisor.  Everything is enqueued back to back; the single read-back at the
    end fetches what the defense reports (trust scores or norms, counts).
    Returns (flat fp32 [d] on the device or None with ``outputs``, info float64 [n], kept, flag)."""
    outs = u.check_outputs(outputs) if outputs is not None else None
    u = u.resident()
    n = u.n
    L = lib()
    with torch.cuda.device(u.device):
        st = _stream()
        g_ptr, has_g0 = None, g0_flat is not None
        if has_g0:
            g0_flat = g0_flat.to(device=u.device, dtype=torch.float32).contiguous()
            if g0_flat.numel() != u.d:
                raise ValueError("g0 has the wrong number of elements")
            g_ptr = g0_flat.data_ptr() or None          # an empty column shard: no pointer, the partial sums stay zero
        # one allocation for every small device array of the round (at C1 / C2 sizes a call is host-bound: each
        # torch.empty / torch.zeros is 5-10 us):  FP64 [2n+1] statistics | FP64 [n+1] info + counts | fp32 [n] linf |
        # fp32 [2n+1] w1, w2, divisor | int32 [n] rows | n bytes keep
        nbytes = 8 * (2 * n + 1) + 8 * (n + 1) + 4 * n + 4 * (2 * n + 1) + 4 * n + n
        scratch = torch.empty(nbytes + 8, dtype=torch.uint8, device=u.device)
        if has_g0 and g_ptr is None:
            scratch.zero_()                             # K2 leaves dot / g0_sumsq unwritten without a g0 pointer; they are all-reduced below
        p = (scratch.data_ptr() + 7) & ~7
        p_info = p + 8 * (2 * n + 1)
        p_linf = p_info + 8 * (n + 1)
        wp = p_linf + 4 * n
        p_rows = wp + 4 * (2 * n + 1)
        p_keep = p_rows + 4 * n
        off = p - scratch.data_ptr()
        buf = scratch[off:off + 8 * (2 * n + 1)].view(torch.float64)
        info = scratch[off + 8 * (2 * n + 1):off + 8 * (3 * n + 2)].view(torch.float64)
        linf = scratch[off + 8 * (3 * n + 2):off + 8 * (3 * n + 2) + 4 * n].view(torch.float32)
        check(L.flagg_row_reduce(u.view(), g_ptr, 0, p, (p + 8 * n) if g_ptr else None, p_linf,
                                 (p + 16 * n) if g_ptr else None, st), "flagg_row_reduce")
        if group is not None:
            all_reduce(buf, group, "sum")
            if use_linf:
                all_reduce(linf, group, "max")
        ne = _examples_on_device(num_examples, u.device) if num_examples is not None else None
        check(L.flagg_round_weights(int(rule), n, None if use_linf else p, (p + 8 * n) if has_g0 else None,
                                    (p + 16 * n) if has_g0 else None, p_linf if use_linf else None,
                                    ne.data_ptr() if ne is not None else None, float(p0), float(p1), p_rows, wp, wp + 4 * n,
                                    wp + 8 * n, p_info, p_info + 8 * n, p_keep, st), "flagg_round_weights")
        out = torch.empty(u.d, dtype=torch.float32, device=u.device) if outs is None else None
        if outs is not None:
            u._bind(outs)
        try:
            check(L.flagg_row_combine_dev(u.view(), n, p_rows, wp, wp + 4 * n, wp + 8 * n,
                                          out.data_ptr() if out is not None else None, None, None, st), "flagg_row_combine_dev")
        finally:
            if outs is not None:
                u._bind(None)
        h = info.cpu()                                                            # the round's one synchronisation
        counts = h[n:].view(torch.int32)
        return out, h[:n].numpy().copy(), int(counts[0]), int(counts[1])


_EXAMPLES = [None, None, None]     # (tuple(num_examples), device, FP64 device tensor) of the latest round


def _examples_on_device(num_examples, device) -> torch.Tensor:
    """The clients' example counts as FP64 on the device.  Rounds of one experiment repeat the same counts
    (the clients' dataset sizes): the upload, a pageable copy, is made when they change."""
    key = tuple(int(x) for x in num_examples)
    if _EXAMPLES[0] != key or _EXAMPLES[1] != device:
        t = torch.tensor([float(x) for x in key], dtype=torch.float64).to(device)
        _EXAMPLES[0], _EXAMPLES[1], _EXAMPLES[2] = key, device, t
    return _EXAMPLES[2]


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
        # up to 31 clients every row is used: the Updates is its own sub-view (no second pointer table / view).
        # Marked with True, not with `u` itself: a self-reference would leave the object (its view, the caller's
        # tensors and a dense device copy of host inputs) to the cyclic GC instead of the refcount.
        sub = u._center_sub = True if len(idx) == u.n else Updates([u.rows[i] for i in idx])
    return colwise_select(u if sub is True else sub, MEDIAN)


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


def krum_scores(G: torch.Tensor, f: int, want_D: bool = True, want_scores: bool = True, squared: bool = False):
    """K5.  (D [n,n] float64 ndarray or None, scores [n] float64 ndarray or None).  ``G``: the centred
    Gram matrix, or with ``squared`` the squared distances of ``sqdist``."""
    n = G.shape[0]
    with torch.cuda.device(G.device):
        # one device buffer and one read-back (one synchronisation) for both results
        nd = n * n if want_D else 0
        buf = torch.empty(nd + (n if want_scores else 0), dtype=torch.float64, device=G.device)
        p = buf.data_ptr()
        fn, name = (lib().flagg_krum_scores_from_sqdist, "flagg_krum_scores_from_sqdist") if squared else \
            (lib().flagg_krum_scores, "flagg_krum_scores")
        check(fn(G.data_ptr(), n, f, p if want_D else None, (p + 8 * nd) if want_scores else None, _stream()), name)
        h = buf.cpu().numpy()
        return (h[:nd].reshape(n, n) if want_D else None), (h[nd:] if want_scores else None)


SMALL_COHORT = 32   # up to here distances come from direct differences (flagg_sqdist): one read of X, no centre


def sqdist(u: Updates, group=None) -> torch.Tensor:
    """K4, small cohorts.  FP64 squared pairwise distances [n, n] on the device by direct differences
    (krum.py:95 subtracts first too); all-reduced over ``group`` when the parameter dimension is sharded
    (partial squared distances add)."""
    u = u.resident()
    if u.n > SMALL_COHORT:
        raise FlaggError(f"sqdist serves cohorts of up to {SMALL_COHORT} clients; larger ones use gram()")
    with torch.cuda.device(u.device):
        D2 = torch.empty((u.n, u.n), dtype=torch.float64, device=u.device)
        check(lib().flagg_sqdist(u.view(), 0, D2.data_ptr(), _stream()), "flagg_sqdist")
        if group is not None:
            all_reduce(D2, group, "sum")
    return D2


def krum_distances_scores(u: Updates, f: int, group=None, want_D: bool = True):
    """Krum's distance matrix and scores (krum.py:86-131): K4 Gram (+ all-reduce when sharded) and
    K5.  A cohort with a non-finite row (a client that sent NaN / Inf) takes the exact path
    below, so that client is ranked last as in the reference instead of poisoning the centre."""
    u = u.resident()
    if u.n <= SMALL_COHORT:
        # direct differences: NaN / Inf rows give NaN / inf distances by IEEE arithmetic, as torch.norm(f_i - f_j) does
        return krum_scores(sqdist(u, group=group), f, want_D, squared=True)
    G = gram(u, group=group)
    D, scores = krum_scores(G, f, want_D)
    if np.all(np.isfinite(scores)):
        return D, scores
    return _krum_nonfinite(u, f, group)


def _krum_nonfinite(u: Updates, f: int, group):
    """Rare path: some row holds NaN / Inf (every score of the Gram path is then non-finite for
    that row, and all of them if it sat among the centre rows).  The reference's distances
    (torch.norm(f_i - f_j), krum.py:95) are NaN / inf for pairs with such a row and ordinary
    numbers for the rest: finite rows go through the Gram of their own sub-view, each non-finite
    row b gets ||x_i - x_b||^2 for all i from one centred K2 pass (fp32 differences, so IEEE
    inf / NaN arithmetic decides inf versus NaN exactly as torch.norm does), K5 ranks the assembled
    matrix with NaN last (np.sort, krum.py:126).  Decisions use all-reduced statistics, so every
    rank of a sharded group takes the same branch."""
    n = u.n
    ss = row_stats(u, group=group)["sumsq"]
    bad = [i for i in range(n) if not np.isfinite(ss[i])]
    good = [i for i in range(n) if np.isfinite(ss[i])]
    D = np.zeros((n, n), dtype=np.float64)
    if len(good) >= 2:
        sub = Updates([u.rows[i] for i in good])
        try:
            Dg, _ = krum_scores(gram(sub, group=group), 0, want_D=True, want_scores=False)
        finally:
            sub.close()
        D[np.ix_(good, good)] = Dg
    with np.errstate(all="ignore"):
        for b in bad:
            xb = torch.cat([t.reshape(-1) for t in u.rows[b]])
            col = np.sqrt(row_stats(u, xb, group=group, centered=True)["sumsq"])
            D[b, :] = col
            D[:, b] = col
    np.fill_diagonal(D, 0.0)
    with torch.cuda.device(u.device):
        Dd = torch.from_numpy(D).to(u.device)
        scores = torch.empty(n, dtype=torch.float64, device=u.device)
        check(lib().flagg_krum_scores_from_distances(Dd.data_ptr(), n, f, scores.data_ptr(), _stream()),
              "flagg_krum_scores_from_distances")
        return D, scores.cpu().numpy()


def weiszfeld(u: Updates, max_iters: int, tolerance: float, group=None):
    """Geometric median by Weiszfeld iterations (trimmed_mean.py:240-262), host out of the loop:
    K1 median start, then per iteration K2 (centred: ||x_i - z||^2, all-reduced when sharded), the
    weights 1/max(dist, 1e-10) and their sum formed on the device, K3 with device-resident weights
    whose epilogue also returns ||z_new - z||^2 (all-reduced when sharded, so every rank stops at
    the same iteration); one 8-byte read-back per iteration decides convergence.
    Returns (flat fp32 [d] on the device, num_iters)."""
    u = u.resident()
    n = u.n
    L = lib()
    with torch.cuda.device(u.device):
        cur = colwise_select(u, MEDIAN)
        new = torch.empty_like(cur)
        ss = torch.empty(n, dtype=torch.float64, device=u.device)
        w = torch.empty(n + 1, dtype=torch.float32, device=u.device)     # weights, then their sum
        delta = torch.empty(1, dtype=torch.float64, device=u.device)
        tol32 = np.float32(tolerance)
        num_iters = max_iters
        for it in range(max_iters):
            st = _stream()
            check(L.flagg_row_reduce(u.view(), cur.data_ptr(), 1, ss.data_ptr(), None, None, None, st), "flagg_row_reduce")
            if group is not None:
                all_reduce(ss, group, "sum")
            check(L.flagg_weiszfeld_weights(ss.data_ptr(), n, w.data_ptr(), w.data_ptr() + 4 * n, st), "flagg_weiszfeld_weights")
            check(L.flagg_row_combine_dev(u.view(), n, None, w.data_ptr(), None, w.data_ptr() + 4 * n, new.data_ptr(),
                                          cur.data_ptr(), delta.data_ptr(), st), "flagg_row_combine_dev")
            if group is not None:
                all_reduce(delta, group, "sum")
            cur, new = new, cur
            with np.errstate(all="ignore"):
                change = np.float32(np.sqrt(delta.cpu().numpy()[0]))
            if change < tol32:                                            # trimmed_mean.py:258-260
                num_iters = it + 1
                break
        return cur, num_iters


def segments_transform(op: int, s: float, x: Sequence[torch.Tensor], y: Optional[Sequence[torch.Tensor]] = None,
                       out: Optional[Sequence[torch.Tensor]] = None) -> List[torch.Tensor]:
    """Elementwise transform of one client's P tensors (flagg_segments_transform): the reference's
    model-poisoning attacks (model_poisoning.py:71-78, 236, 271-283, 172-184) in its own fp32
    operation order.  ``out=None`` allocates fresh tensors; ``out=x`` transforms the rows of a
    resident update matrix in place.  Host tensors are moved to the GPU and the result back."""
    _require_cuda()
    on_host = not x[0].is_cuda
    dev = torch.device("cuda", torch.cuda.current_device()) if on_host else x[0].device

    def prep(ts):
        return [t.detach().to(device=dev, dtype=torch.float32).contiguous() for t in ts]
    xs = prep(x)
    ys = prep(y) if y is not None else None
    if ys is not None and [t.numel() for t in ys] != [t.numel() for t in xs]:
        raise ValueError("transform operands differ in shape")
    if out is None:
        outs = [torch.empty_like(t) for t in xs]
    else:
        outs = list(out)
        for t, o in zip(xs, outs):
            if not (o.is_cuda and o.dtype == torch.float32 and o.is_contiguous() and o.numel() == t.numel()):
                raise ValueError("in-place transforms need contiguous fp32 device tensors of the operands' sizes")
    P = len(xs)
    seg = (C.c_int64 * P)(*[int(t.numel()) for t in xs])
    xp = (C.c_void_p * P)(*[t.data_ptr() for t in xs])
    yp = (C.c_void_p * P)(*[t.data_ptr() for t in ys]) if ys is not None else None
    op_ = (C.c_void_p * P)(*[t.data_ptr() for t in outs])
    with torch.cuda.device(dev):
        check(lib().flagg_segments_transform(P, seg, xp, yp, op_, int(op), float(np.float32(s)), _stream()),
              "flagg_segments_transform")
    if out is None:
        outs = [o.reshape(t.shape) for o, t in zip(outs, x)]
        if on_host:
            outs = [o.cpu() for o in outs]
    return outs


def device_info() -> dict:
    _require_cuda()
    a, b, c = C.c_int(), C.c_int(), C.c_int()
    check(lib().flagg_device_info(C.byref(a), C.byref(b), C.byref(c)), "flagg_device_info")
    return dict(sm_count=a.value, cc=(b.value, c.value))
