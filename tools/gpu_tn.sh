#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_tn_grouped.py tests/test_gpu_dense_tc.py tests/test_gpu_schnet.py tests/test_gpu_painn.py -x -q 2>&1 | tail -5 ) > gpurun_out/pytest_tn.log
cat gpurun_out/pytest_tn.log
for w in schnet painn; do
  b=256; [ $w = painn ] && b=512
  KGCNN_PDL=0 timeout 600 python tools/profile_step.py $w --batch $b --rows 12 > gpurun_out/step_${w}_tn.txt 2>/dev/null
  head -8 gpurun_out/step_${w}_tn.txt | cut -c1-140
  ( timeout 300 python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>/dev/null | tail -1 ) > gpurun_out/bench_${w}_tn.json
  python -c "
import json
d=json.load(open('gpurun_out/bench_${w}_tn.json')); print('$w', 'ms/step %.3f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'])"
done
