"""pytest configuration: markers, import paths, golden-fixture loader."""
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "multimodal-fl-security_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


class Golden:
    """tests/golden/golden.npz + golden.json, written by tests/golden/make_golden.py
    from the reference's own classes."""

    def __init__(self):
        g = os.path.join(ROOT, "tests", "golden")
        with open(os.path.join(g, "golden.json")) as fh:
            self.meta = json.load(fh)
        self.arr = np.load(os.path.join(g, "golden.npz"))

    def names(self, prefix=""):
        return sorted(k for k in self.meta["cases"] if k.startswith(prefix))

    def case(self, name):
        c = dict(self.meta["cases"][name])
        c["X"] = self.arr[f"{name}/X"]
        for f in c["fields"]:
            c[f] = self.arr[f"{name}/{f}"]
        return c


_GOLDEN = Golden()


@pytest.fixture(scope="session")
def golden():
    return _GOLDEN


def golden_names(prefix):
    return _GOLDEN.names(prefix)


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def same_bits_nan_aware(got, ref):
    """Bit-identical except that a NaN may carry any payload: x86 and the GPU propagate / generate
    different NaN bit patterns through the same arithmetic (0x7FC00000 vs 0x7FFFFFFF); position and
    NaN-ness must agree, every other value bit for bit."""
    got = np.ascontiguousarray(got, dtype=np.float32)
    ref = np.ascontiguousarray(ref, dtype=np.float32)
    gn, rn = np.isnan(got), np.isnan(ref)
    return got.shape == ref.shape and np.array_equal(gn, rn) and np.array_equal(bits(got)[~gn], bits(ref)[~rn])


def assert_close_scaled(got, ref, scale, tol=1e-5, what=""):
    """|got-ref| <= tol*(|ref|+scale) per coordinate and ||got-ref|| <= tol*||ref||
    (SURVEY.md section 7 'Tolerance definition'); NaN/Inf positions must agree exactly."""
    got = np.asarray(got, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    fin = np.isfinite(ref)
    assert np.array_equal(np.isnan(got), np.isnan(ref)), f"{what}: NaN positions differ"
    assert np.array_equal(got[~fin & ~np.isnan(ref)], ref[~fin & ~np.isnan(ref)]), f"{what}: Inf differ"
    err = np.abs(got[fin] - ref[fin])
    bound = tol * (np.abs(ref[fin]) + np.broadcast_to(scale, ref.shape)[fin])
    worst = (err / np.maximum(bound, 1e-300)).max() if err.size else 0.0
    assert worst <= 1.0, f"{what}: element-wise error {worst:.3g}x the 1e-5 scaled bound"
    nr = np.linalg.norm(ref[fin])
    if nr > 0:
        assert np.linalg.norm(got[fin] - ref[fin]) <= tol * nr, f"{what}: norm-wise error too large"
