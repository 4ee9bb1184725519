#!/bin/bash
# K4 development visit: TMEM layout probe, Gram check (ts / ss / ffma) with timings.
mkdir -p gpurun_out
nvcc -gencode arch=compute_100a,code=sm_100a -o /tmp/tmem_probe tools/tmem_probe.cu && timeout 60 /tmp/tmem_probe > gpurun_out/tmem_probe.log 2>&1; echo "probe rc=$?" >> gpurun_out/tmem_probe.log
cat gpurun_out/tmem_probe.log
timeout 900 python tools/gram_tc_check.py --peak --time > gpurun_out/gram_check.log 2>&1; echo "check rc=$?" >> gpurun_out/gram_check.log
cat gpurun_out/gram_check.log
