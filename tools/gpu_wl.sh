#!/bin/bash
# one workload: bench line + per-kernel step table.  usage: gpu_wl.sh <schnet|painn|hdnnp> [batch]
mkdir -p gpurun_out
w=$1; b=${2:-256}
( timeout 600 python bench.py --steps 5 --warmup 3 --workload $w --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_$w.err | tail -2 ) > gpurun_out/bench_$w.json
( timeout 600 python tools/profile_step.py $w --batch $b 2>&1 | tail -60 ) > gpurun_out/step_$w.txt
cut -c1-300 gpurun_out/bench_$w.json; tail -5 gpurun_out/bench_$w.err; cat gpurun_out/step_$w.txt
