#!/usr/bin/env python
"""K4 timing for n <= 256 under each tensor-core kernel (GPU box only)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
from gram_tc_check import timing
for n, d in ((64, 6_250_000), (100, 6_250_000), (128, 6_250_000), (200, 3_125_000), (256, 3_125_000)):
    for impl in sys.argv[1:] or ["ss", "ts1", "ts2"]:
        timing(n, d, impl, reps=5)
