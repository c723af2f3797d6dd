#!/bin/bash
# fused cfconv: kernel-level parity, then the SchNet model tests and a short C2 bench
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_cfconv.py -x -q 2>&1 | tail -25 ) > gpurun_out/pytest_cf.log
cat gpurun_out/pytest_cf.log
( timeout 900 python -m pytest tests/test_gpu_schnet.py tests/test_gpu_parity_full.py tests/test_gpu_capacity.py tests/test_gpu_fullsize.py -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_schnet.log
cat gpurun_out/pytest_schnet.log
( timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_schnet.err | tail -2 ) > gpurun_out/bench_schnet.json
cut -c1-300 gpurun_out/bench_schnet.json; tail -3 gpurun_out/bench_schnet.err
timeout 600 python tools/profile_step.py schnet --batch 256 --rows 40 > gpurun_out/step_schnet_cf.txt 2>gpurun_out/step_schnet_cf.err
head -30 gpurun_out/step_schnet_cf.txt | cut -c1-150
