#!/bin/bash
# n <= 128 Gram: timing of ss / ts1, then one ncu --set full capture of the ts1 kernel at n = 128 x 6.25M
mkdir -p gpurun_out
timeout 300 python tools/k4_small_n.py ss ts1 > gpurun_out/k4_small_n_ts1.txt 2>&1
cat gpurun_out/k4_small_n_ts1.txt
K4_N=128 K4_D=6250000 timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram_ts1_kernel -s 1 -c 1 -o gpurun_out/k4_ts1 -f python tools/gram_tc_check.py --only-time ts1 > gpurun_out/ncu_k4_ts1.log 2>&1
K4_N=128 K4_D=6250000 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/k4_ts1_launches.csv python tools/gram_tc_check.py --only-time ts1 > /dev/null 2>&1
tail -3 gpurun_out/ncu_k4_ts1.log
