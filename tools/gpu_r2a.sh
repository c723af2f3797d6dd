#!/bin/bash
# round-2 pass A: parity tests (incl. the full-config fp64 comparisons), smoke, default bench
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/pytest_gpu.log
( timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -5 ) > gpurun_out/smoke.log
( timeout 600 python bench.py --steps 10 --warmup 3 2>gpurun_out/bench_schnet.err | tail -2 ) > gpurun_out/bench_schnet.json
tail -25 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cut -c1-600 gpurun_out/bench_schnet.json; tail -3 gpurun_out/bench_schnet.err
