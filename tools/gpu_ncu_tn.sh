#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_step.py schnet > gpurun_out/ncu_tn_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:tn_grouped_kernel -c 2 -f -o gpurun_out/tn_grouped_c2 python tools/ncu_step.py schnet > gpurun_out/ncu_tn.log 2>&1
tail -2 gpurun_out/ncu_tn.log
ncu -i gpurun_out/tn_grouped_c2.ncu-rep --page raw --csv > gpurun_out/r2_tn_grouped_c2_raw.csv 2>/dev/null
ncu -i gpurun_out/tn_grouped_c2.ncu-rep --page source --csv --print-source sass > gpurun_out/r2_tn_grouped_c2_source.csv 2>/dev/null
ls -la gpurun_out | grep tn_
