#!/usr/bin/env python
"""K4 timing for n < 64: CUDA-core kernel against the single-CTA tensor kernel (GPU box only)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
from gram_tc_check import timing
for n, d in ((10, 11_689_512), (20, 11_689_512), (32, 11_689_512), (50, 11_689_512), (63, 6_250_000)):
    for impl in sys.argv[1:] or ["ts1", "small"]:
        timing(n, d, impl, reps=5)
