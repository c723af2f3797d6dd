#!/bin/bash
# dense kernel iteration: correctness tests, timeline probe, microbench
mkdir -p gpurun_out
( timeout 600 python -m pytest tests -m gpu -x -q -k "dense_tc or tn_grouped or weight_gradient_tn" 2>&1 | tail -8 ) > gpurun_out/pytest_quick.log
( timeout 120 tools/tc_timeline 18432 2>&1 ) > gpurun_out/timeline.txt
( timeout 600 python tools/mb_dense.py 2>&1 | tail -20 ) > gpurun_out/mb_dense.txt
tail -8 gpurun_out/pytest_quick.log; head -21 gpurun_out/timeline.txt; grep -A17 "NN+act" gpurun_out/timeline.txt | tail -6;  cat gpurun_out/mb_dense.txt
