#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_tn_grouped.py tests/test_gpu_dense_tc.py tests/test_gpu_schnet.py tests/test_gpu_painn.py -x -q 2>&1 | tail -5 ) > gpurun_out/pytest_tn.log
cat gpurun_out/pytest_tn.log
for w in schnet painn; do
  b=256; [ $w = painn ] && b=512
  timeout 600 python tools/profile_step.py $w --batch $b --rows 12 > gpurun_out/step_${w}_tn.txt 2>/dev/null
  head -8 gpurun_out/step_${w}_tn.txt | cut -c1-140
done
