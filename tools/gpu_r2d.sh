#!/bin/bash
# ncu capture of the roofline kernels + the default bench line (incl. other_workloads) + reference arm
mkdir -p gpurun_out
python tools/ncu_gms.py > gpurun_out/plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:"gms_kernel|gather2_mul" -f -o gpurun_out/gms_full python tools/ncu_gms.py > gpurun_out/ncu_gms.log 2>&1
tail -3 gpurun_out/ncu_gms.log
( timeout 900 python bench.py 2>gpurun_out/bench_default.err | tail -1 ) > gpurun_out/bench_default.json
cut -c1-1500 gpurun_out/bench_default.json; tail -3 gpurun_out/bench_default.err
