#!/bin/bash
# One GPU-box pass: parity tests, smoke, every bench workload; logs under gpurun_out/.
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log
( timeout 120 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -5 ) > gpurun_out/smoke.log
for w in schnet painn hdnnp gcn gss; do
  ( timeout 600 python bench.py --workload $w --steps 10 --warmup 3 2>gpurun_out/bench_$w.err | tail -2 ) > gpurun_out/bench_$w.json
done
( timeout 300 python bench.py --impl reference --steps 2 --warmup 1 2>&1 | tail -2 ) > gpurun_out/bench_ref.json
tail -3 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; cat gpurun_out/bench_*.json
