#!/bin/bash
# ncu --set full capture of the K4 kernel selected by $1 (regex) under FLAGG_GRAM_IMPL=$2
mkdir -p gpurun_out
FLAGG_GRAM_IMPL_FORCE=$2 timeout 600 ncu --set full --clock-control none --import-source on -k regex:$1 -s 1 -c 1 -o gpurun_out/k4_$2 -f python tools/gram_tc_check.py --only-time $2 > gpurun_out/ncu_k4_$2.log 2>&1
tail -3 gpurun_out/ncu_k4_$2.log
ls -la gpurun_out
