#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q -k "dense or primitives or linear or parity or relational" 2>&1 | tail -6 ) > gpurun_out/pytest_quick.log
( timeout 600 python tools/mb_dense.py 2>&1 | tail -20 ) > gpurun_out/mb_dense.txt
( timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_schnet.err | tail -2 ) > gpurun_out/bench_schnet.json
cat gpurun_out/pytest_quick.log gpurun_out/mb_dense.txt; cut -c1-260 gpurun_out/bench_schnet.json
