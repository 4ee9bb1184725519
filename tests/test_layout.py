"""CPU checks of the boundary: libflagg.so loads and exports every symbol include/flagg.h
declares; the product tree never touches oracle/ and has no CPU fallback."""
import os
import re

import pytest
import torch

from conftest import PKG, ROOT


def header_symbols():
    src = open(os.path.join(ROOT, "include", "flagg.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(flagg_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import ctypes
    import flagg
    lib = flagg.lib()
    names = header_symbols()
    assert len(names) >= 14
    for s in names:
        assert hasattr(lib, s), f"{s} declared in include/flagg.h but not exported"
    assert sorted(flagg.EXPORTS) == names
    assert lib.flagg_version() == 100
    assert isinstance(lib.flagg_last_error(), bytes)
    assert ctypes.sizeof(ctypes.c_void_p) == 8


def test_product_never_imports_oracle():
    bad = []
    for dp, _, files in os.walk(PKG):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                txt = open(os.path.join(dp, f)).read()
                if re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M) or "oracle/" in txt:
                    bad.append(os.path.join(dp, f))
    assert not bad, bad


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without a GPU")
def test_no_cpu_fallback_without_gpu():
    from flagg import FlaggError
    from src.defenses import get_defense
    ups = [[torch.zeros(8)] for _ in range(5)]
    for name, cfg in [("median", {}), ("trimmed_mean", {}), ("krum", {"num_malicious": 1}), ("fedavg", {}),
                      ("gradient_clipping", {}), ("norm_bounding", {}), ("dp_sgd", {})]:
        with pytest.raises(FlaggError):
            get_defense(name, cfg).aggregate(ups, [1] * 5)


def test_factory_names_and_errors():
    from src.defenses import get_defense
    for name in ["none", "fedavg", "krum", "multi_krum", "trimmed_mean", "median", "geometric_median", "fltrust",
                 "dp_sgd", "gradient_clipping", "norm_bounding"]:
        assert get_defense(name, {}) is not None
    with pytest.raises(ValueError):
        get_defense("bogus", {})
    with pytest.raises(ValueError):      # validation happens before any kernel is needed
        get_defense("krum", {"num_malicious": 2}).aggregate([[torch.zeros(2)]] * 3, [1] * 3)
    with pytest.raises(ValueError):
        get_defense("fltrust", {}).aggregate([[torch.zeros(2)]], [1])
