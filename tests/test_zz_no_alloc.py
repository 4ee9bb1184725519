"""GPU test of flagg_row_reduce_static, the K2 entry point that allocates nothing (include/flagg.h: its
per-CTA partial sums live in a file-scope ``static __device__`` array of FLAGG_STATIC_SCRATCH_BYTES).
It has no counterpart in the reference; the checker is the existing entry point, flagg_row_reduce, on the
same view: same kernels, same grid, hence the same bits.  Called straight through the C-ABI with ctypes.

Collected last on purpose (``pytest -x`` stops at the first failure) and self-contained: it passes alone in
a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

EUNSUPPORTED = -4


def _lib():
    from flagg import _lib as binding
    return binding, binding.lib()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class View:
    """flagg_view over n rows x P separately held tensors (torch only owns the memory)."""

    def __init__(self, rows, seg):
        binding, L = _lib()
        self.rows, self.n, self.P = rows, len(rows), len(seg)
        seg_arr = (C.c_int64 * self.P)(*seg)
        ptrs = (C.c_void_p * (self.n * self.P))(*[t.data_ptr() for row in rows for t in row])
        self.h = C.c_void_p()
        binding.check(L.flagg_view_create(C.byref(self.h), self.n, self.P, seg_arr, ptrs, _stream()), "flagg_view_create")

    def close(self):
        binding, L = _lib()
        torch.cuda.synchronize()
        binding.check(L.flagg_view_destroy(self.h), "flagg_view_destroy")


def _make(n, seg, seed, offset=0):
    """Rows of fp32 tensors; offset > 0 shifts every base by 4 * offset bytes (the scalar-load kernels)."""
    g = torch.Generator(device="cuda").manual_seed(seed)
    rows = []
    for i in range(n):
        scale = 1e-3 * (10.0 if i < 2 else 1.0)
        rows.append([(torch.randn(s + offset, device="cuda", generator=g) * scale + 0.05)[offset:] for s in seg])
    d = sum(seg)
    g0 = torch.randn(d + offset, device="cuda", generator=g)[offset:] * 1e-3
    return rows, g0


def _outputs(n):
    # poisoned, so that an output the call did not write shows up
    return dict(sumsq=torch.full((n,), -7.0, dtype=torch.float64, device="cuda"),
                dot=torch.full((n,), -7.0, dtype=torch.float64, device="cuda"),
                linf=torch.full((n,), -7.0, dtype=torch.float32, device="cuda"),
                g0ss=torch.full((1,), -7.0, dtype=torch.float64, device="cuda"))


def _call(fn_name, view, g0, centered, with_g0_stats=True):
    binding, L = _lib()
    o = _outputs(view.n)
    has_g0 = g0 is not None
    rc = getattr(L, fn_name)(view.h, g0.data_ptr() if has_g0 else None, int(centered), o["sumsq"].data_ptr(),
                             o["dot"].data_ptr() if has_g0 else None, o["linf"].data_ptr(),
                             o["g0ss"].data_ptr() if has_g0 and with_g0_stats else None, _stream())
    binding.check(rc, fn_name)
    torch.cuda.synchronize()
    return o, has_g0


def _same(a, b, has_g0, what):
    keys = ["sumsq", "linf"] + (["dot", "g0ss"] if has_g0 else [])
    for k in keys:
        x, y = a[k].cpu().numpy(), b[k].cpu().numpy()
        assert np.array_equal(x.view(np.uint8), y.view(np.uint8)), f"{what}: {k} differs from flagg_row_reduce: {x[:4]} vs {y[:4]}"
        assert not np.any(x == -7.0), f"{what}: {k} was not written"


def _check_against_fp64(view, o, g0, centered):
    """The pair also agrees with plain FP64 arithmetic (1e-5, north_star's tolerance; fp32 products inside)."""
    X = torch.stack([torch.cat([t.reshape(-1) for t in row]) for row in view.rows]).double()
    if centered:
        X = X - g0.double()
    np.testing.assert_allclose(o["sumsq"].cpu().numpy(), (X * X).sum(dim=1).cpu().numpy(), rtol=1e-5)
    np.testing.assert_allclose(o["linf"].cpu().numpy(), X.abs().max(dim=1)[0].float().cpu().numpy(), rtol=1e-6)


CASES = [
    # n, segments, element offset of every base, g0?, centered
    (20, [5000, 77, 12345], 0, True, 0),
    (20, [5000, 77, 12345], 0, True, 1),
    (50, [40_000, 333, 60_000], 0, False, 0),      # no g0: the vectorised kernel on staged zeros
    (13, [4099, 1, 2048], 1, True, 0),             # 4-byte aligned bases: scalar loads
    (13, [4099, 1, 2048], 1, False, 0),
    (500, [150_000, 50_000], 0, True, 0),          # C5's cohort size
]


@pytest.mark.parametrize("n,seg,offset,use_g0,centered", CASES)
def test_static_scratch_matches_row_reduce_two_calls_in_a_row(n, seg, offset, use_g0, centered):
    rows, g0 = _make(n, seg, 1000 + n + centered)
    if offset:
        rows, g0 = _make(n, seg, 1000 + n, offset=offset)
    view = View(rows, seg)
    # a second, different and smaller problem: the shared array still holds the first call's partials when it runs
    rows2, g02 = _make(max(3, n // 2), [seg[0] // 2 + 1, 513], 2000 + n, offset=offset)
    view2 = View(rows2, [seg[0] // 2 + 1, 513])
    try:
        ref1, has = _call("flagg_row_reduce", view, g0 if use_g0 else None, centered)
        ref2, _ = _call("flagg_row_reduce", view2, g02 if use_g0 else None, centered)
        got1, _ = _call("flagg_row_reduce_static", view, g0 if use_g0 else None, centered)
        got2, _ = _call("flagg_row_reduce_static", view2, g02 if use_g0 else None, centered)
        got1b, _ = _call("flagg_row_reduce_static", view, g0 if use_g0 else None, centered)
        _same(got1, ref1, has, "first call")
        _same(got2, ref2, has, "second call (stale partials in the array)")
        _same(got1b, ref1, has, "third call, first problem again")
        _check_against_fp64(view, got1, g0, centered)
        _check_against_fp64(view2, got2, g02, centered)
    finally:
        view.close()
        view2.close()


def test_static_scratch_back_to_back_without_a_host_sync():
    """Two calls enqueued on one stream with nothing between them: stream order is the only ordering the shared
    array needs."""
    binding, L = _lib()
    seg = [30_000, 1025]
    rows_a, g0_a = _make(24, seg, 5)
    rows_b, g0_b = _make(24, seg, 6)
    va, vb = View(rows_a, seg), View(rows_b, seg)
    try:
        ref_a, _ = _call("flagg_row_reduce", va, g0_a, 0)
        ref_b, _ = _call("flagg_row_reduce", vb, g0_b, 0)
        oa, ob = _outputs(24), _outputs(24)
        for v, g, o in ((va, g0_a, oa), (vb, g0_b, ob)):
            binding.check(L.flagg_row_reduce_static(v.h, g.data_ptr(), 0, o["sumsq"].data_ptr(), o["dot"].data_ptr(),
                                                    o["linf"].data_ptr(), o["g0ss"].data_ptr(), _stream()), "flagg_row_reduce_static")
        torch.cuda.synchronize()
        _same(oa, ref_a, True, "first of two queued calls")
        _same(ob, ref_b, True, "second of two queued calls")
    finally:
        va.close()
        vb.close()


def test_static_scratch_takes_no_device_memory():
    """Device-wide free memory is the same before and after a call (the array is part of the loaded module)."""
    seg = [1 << 20]
    rows, g0 = _make(64, seg, 9)
    view = View(rows, seg)
    try:
        _call("flagg_row_reduce_static", view, g0, 0)        # first use on this device: symbol lookup, kernel attributes
        o = _outputs(view.n)
        binding, L = _lib()
        torch.cuda.synchronize()
        free0, _ = torch.cuda.mem_get_info()
        for _ in range(3):
            binding.check(L.flagg_row_reduce_static(view.h, g0.data_ptr(), 0, o["sumsq"].data_ptr(), o["dot"].data_ptr(),
                                                    o["linf"].data_ptr(), o["g0ss"].data_ptr(), _stream()), "flagg_row_reduce_static")
        torch.cuda.synchronize()
        free1, _ = torch.cuda.mem_get_info()
        assert free1 == free0, f"free device memory changed by {free0 - free1} bytes across flagg_row_reduce_static"
    finally:
        view.close()


def test_static_scratch_rejects_sizes_beyond_the_array():
    """20 n + 8 bytes of partials per CTA, 4 x SM count CTAs once d >= 2048 x that: on 148 SMs n = 708 is the
    last cohort that fits 8 MiB.  The rows alias one tensor (the view only holds pointers), so this is cheap."""
    binding, L = _lib()
    sms = torch.cuda.get_device_properties(0).multi_processor_count
    ctas = 4 * sms
    d = 2048 * ctas
    n_fit = (binding.STATIC_SCRATCH_BYTES // ctas - 8) // 20
    shared = torch.randn(d, device="cuda") * 1e-3
    for n, fits in ((n_fit, True), (n_fit + 1, False)):
        view = View([[shared]] * n, [d])
        try:
            o = _outputs(n)
            rc = L.flagg_row_reduce_static(view.h, None, 0, o["sumsq"].data_ptr(), None, o["linf"].data_ptr(), None, _stream())
            torch.cuda.synchronize()
            if fits:
                binding.check(rc, "flagg_row_reduce_static")
                ref, _ = _call("flagg_row_reduce", view, None, 0)
                _same(o, ref, False, f"n = {n}, the largest cohort that fits")
            else:
                assert rc == EUNSUPPORTED
                msg = L.flagg_last_error().decode()
                assert "FLAGG_STATIC_SCRATCH_BYTES" in msg and "flagg_row_reduce" in msg, msg
                assert torch.all(o["sumsq"] == -7.0), "a rejected call must not launch anything"
        finally:
            view.close()


def test_static_scratch_argument_errors_and_empty_view():
    binding, L = _lib()
    assert L.flagg_row_reduce_static(None, None, 0, None, None, None, None, None) == -1
    rows, g0 = _make(4, [100], 3)
    view = View(rows, [100])
    try:
        o = _outputs(4)
        # dot without g0 is an argument error, as in flagg_row_reduce
        assert L.flagg_row_reduce_static(view.h, None, 0, o["sumsq"].data_ptr(), o["dot"].data_ptr(), None, None, _stream()) == -1
        assert L.flagg_row_reduce_static(view.h, None, 1, o["sumsq"].data_ptr(), None, None, None, _stream()) == -1
    finally:
        view.close()
    # an empty column shard (d = 0): zeros, as flagg_row_reduce gives
    empty = View([[torch.empty(0, device="cuda")] for _ in range(4)], [0])
    try:
        got, _ = _call("flagg_row_reduce_static", empty, None, 0)
        ref, _ = _call("flagg_row_reduce", empty, None, 0)
        assert torch.all(got["sumsq"] == 0) and torch.all(got["linf"] == 0)
        assert torch.equal(got["sumsq"], ref["sumsq"]) and torch.equal(got["linf"], ref["linf"])
    finally:
        empty.close()
