#!/bin/bash
# ncu --set full of K1 on the gather4 feed (256-column tiles) at C3, after the command exited 0 without ncu
mkdir -p gpurun_out
timeout 300 python tools/k1_layouts.py 100 25000000 > gpurun_out/k1_layouts.txt 2>&1 || exit 1
timeout 200 python tools/k1_layouts.py 20 11689512 trimmed >> gpurun_out/k1_layouts.txt 2>&1
timeout 200 python tools/k1_layouts.py 64 10000000 >> gpurun_out/k1_layouts.txt 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:colselect_bulk_kernel -s 30 -c 1 -o gpurun_out/k1_gather_full -f python tools/k1_layouts.py 100 25000000 > gpurun_out/ncu_k1_gather.log 2>&1
tail -2 gpurun_out/ncu_k1_gather.log
