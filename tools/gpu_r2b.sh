#!/bin/bash
# dense microbench + full-config parity + bench
mkdir -p gpurun_out
( timeout 600 python tools/mb_dense.py 2>&1 | tail -20 ) > gpurun_out/mb_dense.txt
( timeout 900 python -m pytest tests -m gpu -x -q -k "parity_full" 2>&1 | tail -15 ) > gpurun_out/pytest_quick.log
( timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_schnet.err | tail -2 ) > gpurun_out/bench_schnet.json
( timeout 600 python bench.py --steps 5 --warmup 3 --workload painn --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_painn.err | tail -2 ) > gpurun_out/bench_painn.json
cat gpurun_out/mb_dense.txt; tail -15 gpurun_out/pytest_quick.log; cut -c1-330 gpurun_out/bench_schnet.json; cut -c1-330 gpurun_out/bench_painn.json
