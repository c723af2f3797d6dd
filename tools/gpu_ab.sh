#!/bin/bash
# A/B of one environment switch on the C2 step inside ONE box: bash tools/gpu_ab.sh VAR
VAR=${1:-KGCNN_CFCONV_MERGE_WG}
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_cfconv.py -x -q 2>&1 | tail -2 )
for rep in 1 2; do for v in 1 0; do
  ( env $VAR=$v timeout 300 python bench.py --workload schnet --steps 30 --warmup 5 --no-cpu-baseline --no-roofline --no-other-workloads 2>/dev/null | tail -1 ) > gpurun_out/ab_$v.json
  python -c "
import json
d=json.load(open('gpurun_out/ab_$v.json')); print('$VAR=$v', 'ms/step %.3f' % d['ms_per_step'], 'e2e %.0f' % d['e2e']['value'])"
done; done
