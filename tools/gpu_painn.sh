#!/bin/bash
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q -k "dense or painn or primitives or linear" 2>&1 | tail -8 ) > gpurun_out/pytest_quick.log
( timeout 600 python tools/mb_dense.py 2>&1 | tail -20 ) > gpurun_out/mb_dense.txt
bash tools/gpu_wl.sh painn 512 > gpurun_out/wl_painn.txt 2>&1
cat gpurun_out/pytest_quick.log gpurun_out/mb_dense.txt; head -50 gpurun_out/wl_painn.txt
