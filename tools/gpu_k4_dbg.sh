#!/bin/bash
# tuning builds of the A-from-TMEM Gram kernels (libflagg_<variant>.so, see csrc/Makefile): n = 64 / 128 single CTA, n = 256 / 500 / 1000 pairs
for lib in multimodal-fl-security_b200/flagg/libflagg_*.so; do
  echo "== $lib"
  for n in 64 128; do
    FLAGG_LIB=$PWD/$lib K4_N=$n K4_D=6250000 timeout 120 python tools/gram_tc_check.py --only-time ts1 2>&1 | tail -1
  done
  for n in 256 500 1000; do
  FLAGG_LIB=$PWD/$lib K4_N=$n timeout 120 python tools/gram_tc_check.py --only-time ts2 2>&1 | tail -1
  done
done
