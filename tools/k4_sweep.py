#!/usr/bin/env python
"""K4 timing sweep over d (page-count / latency experiment) -- GPU box only."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tools")]
from gram_tc_check import timing
for impl in sys.argv[1:] or ["tc", "ss"]:
    for n, d, reps in ((1000, 100_000, 20), (1000, 400_000, 10), (1000, 1_000_000, 5), (1000, 3_125_000, 3), (512, 3_125_000, 3), (256, 6_250_000, 3)):
        timing(n, d, impl, reps=reps)
