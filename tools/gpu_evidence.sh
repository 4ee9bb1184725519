#!/bin/bash
# Round-2 evidence on one GPU: ncu --set full of K2 / K3 at C5 and of K1 (dense + gather4 feeds) at C3,
# the launch list of the bench command, SASS inventory.  Run only after the same commands exited 0 without ncu.
mkdir -p gpurun_out
timeout 600 python tools/c5_once.py > gpurun_out/c5_once.log 2>&1 || exit 1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'row_reduce_kernel|row_combine_kernel' -s 4 -c 2 -o gpurun_out/k2k3_full -f python tools/c5_once.py > gpurun_out/ncu_k2k3.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:colselect_bulk -s 6 -c 1 -o gpurun_out/k1_gather_full -f python tools/k1_layouts.py 100 25000000 > gpurun_out/ncu_k1_gather.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:colselect_bulk -s 3 -c 1 -o gpurun_out/k1_full -f python bench.py --steps 3 --warmup 3 --no-extras --no-cpu --no-e2e > gpurun_out/ncu_k1.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-extras --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1
ls -la gpurun_out/*.ncu-rep
