#!/bin/bash
mkdir -p gpurun_out
python tools/ncu_cfconv.py > gpurun_out/ncu_cf_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:cfconv_kernel -s 2 -c 2 -f -o gpurun_out/cfconv_full python tools/ncu_cfconv.py > gpurun_out/ncu_cf.log 2>&1
tail -3 gpurun_out/ncu_cf.log
ncu -i gpurun_out/cfconv_full.ncu-rep --page raw --csv > gpurun_out/r2_cfconv_full.csv 2>/dev/null
ls -la gpurun_out | tail -5
