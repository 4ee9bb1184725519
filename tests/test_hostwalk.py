"""The native walk over the plugin API's List[List[Tensor]] (csrc/hostwalk.cpp) against the general Python
constructor: same pointer table, same metadata, and every input the fast path cannot take falls through to the
Python loop, which makes the conversion or words the error.  CPU only (host tensors)."""
import ctypes as C

import pytest
import torch


def _both(ups):
    from flagg import engine
    assert engine._hostwalk() is not None, "flagg/_hostwalk.so was not built (make -C multimodal-fl-security_b200/csrc)"
    fast = engine.Updates(ups)
    saved = engine._HOSTWALK
    engine._HOSTWALK = None
    try:
        slow = engine.Updates(ups)
    finally:
        engine._HOSTWALK = saved
    return fast, slow


def test_fast_walk_equals_python_walk():
    ups = [[torch.randn(3, 5), torch.randn(7), torch.randn(2, 2, 2)] for _ in range(17)]
    fast, slow = _both(ups)
    assert fast._ptrs is not None and slow._ptrs is None
    for a in ("n", "P", "shapes", "numel", "in_dtypes", "d", "on_host"):
        assert getattr(fast, a) == getattr(slow, a), a
    (_, pf), (_, ps) = fast._ptr_tables(), slow._ptr_tables()
    assert list(pf) == list(ps) == [t.data_ptr() for u in ups for t in u]
    assert fast.rows is ups                                  # the caller's lists keep the tensors alive


@pytest.mark.parametrize("what", ["dtype", "noncontig", "grad", "tuple_rows"])
def test_general_path_cases(what):
    ups = [[torch.randn(4, 6), torch.randn(5)] for _ in range(6)]
    if what == "dtype":
        ups[2][1] = ups[2][1].double()
    elif what == "noncontig":
        ups[3][0] = torch.randn(6, 4).t()
    elif what == "grad":
        ups[1][0] = ups[1][0].requires_grad_()
    elif what == "tuple_rows":
        ups = tuple(tuple(u) for u in ups)
    fast, slow = _both(ups)
    assert (fast._ptrs is not None) == (what == "tuple_rows")
    for a in ("n", "P", "shapes", "numel", "d", "on_host"):
        assert getattr(fast, a) == getattr(slow, a), a
    for rf, rs in zip(fast.rows, slow.rows):
        for tf, ts in zip(rf, rs):
            assert tf.dtype == torch.float32 and tf.is_contiguous() and not tf.requires_grad and torch.equal(tf, ts)


def test_errors_are_worded_by_the_python_path():
    from flagg import engine
    ups = [[torch.randn(4), torch.randn(5)] for _ in range(4)]
    ups[2] = [torch.randn(4)]
    with pytest.raises(ValueError, match="client 2 has 1 tensors"):
        engine.Updates(ups)
    ups[2] = [torch.randn(4), torch.randn(6)]
    with pytest.raises(ValueError, match="client 2 tensor 1"):
        engine.Updates(ups)
    with pytest.raises(ValueError):
        engine.Updates([])
