#!/bin/bash
# PDL A/B: parity suite, then the C2 / C3 / C4 steps with and without the launch attribute
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 ) > gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for pdl in 1 0; do
  for w in schnet painn hdnnp; do
    ( KGCNN_PDL=$pdl timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_${w}_pdl$pdl.err | tail -1 ) > gpurun_out/bench_${w}_pdl$pdl.json
    python -c "
import json,sys
d=json.load(open('gpurun_out/bench_${w}_pdl$pdl.json')); print('PDL=$pdl', '$w', 'ms/step %.3f' % d['ms_per_step'], 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'])" || tail -3 gpurun_out/bench_${w}_pdl$pdl.err
  done
done
