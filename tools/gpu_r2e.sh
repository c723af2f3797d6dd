#!/bin/bash
# Re-entry pass: full GPU parity suite, smoke, step tables (CUPTI) for the three force workloads, default bench line.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log
( timeout 200 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -5 ) > gpurun_out/smoke.log
for w in schnet painn hdnnp; do
  b=256; [ $w = painn ] && b=512
  timeout 600 python tools/profile_step.py $w --batch $b --rows 70 > gpurun_out/step_$w.txt 2>gpurun_out/step_$w.err
done
( timeout 900 python bench.py 2>gpurun_out/bench_default.err | tail -1 ) > gpurun_out/bench_default.json
tail -6 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log; head -5 gpurun_out/step_*.txt; cut -c1-1200 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
