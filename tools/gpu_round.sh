#!/bin/bash
# One GPU-box visit: parity tests, smoke, bench, ncu launch list, ncu --set full captures of K1 and K4, small-n K4 timings.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,memory.total --format=csv > gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py --smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv \
    python bench.py --steps 3 --warmup 3 --no-extras --no-cpu --no-e2e > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:colselect_ -s 3 -c 1 -o gpurun_out/k1_full -f \
    python bench.py --steps 3 --warmup 3 --no-extras --no-cpu --no-e2e > gpurun_out/ncu_full.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram_ts2_kernel -s 1 -c 1 -o gpurun_out/k4_full -f python tools/gram_tc_check.py --only-time > gpurun_out/ncu_k4.log 2>&1
K4_N=20 K4_D=11689512 timeout 600 ncu --set full --clock-control none --import-source on -k regex:gram_small_kernel -s 1 -c 1 -o gpurun_out/k4_small -f python tools/gram_tc_check.py --only-time small > gpurun_out/ncu_k4_small.log 2>&1
timeout 300 python tools/k4_tiny_n.py small ts1 2>&1 | grep impl > gpurun_out/k4_tiny_n.txt
timeout 300 python tools/k4_small_n.py ss ts1 2>&1 | grep impl > gpurun_out/k4_small_n_ts1.txt
timeout 300 python tools/api_latency.py > gpurun_out/api_latency.log 2>&1
ls -la gpurun_out
