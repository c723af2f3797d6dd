#!/bin/bash
# Profiling pass (diagnostics, not bench numbers): CUPTI step tables, ncu launch list of the default bench command,
# one ncu --set full capture of the north-star kernel.
mkdir -p gpurun_out
for w in schnet painn hdnnp; do
  b=256; [ $w = painn ] && b=512
  timeout 600 python tools/profile_step.py $w --batch $b --rows 60 > gpurun_out/step_$w.txt 2>/dev/null
done
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/b.log 2>&1 && \
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 12000 --csv --log-file gpurun_out/launches_schnet.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launch.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:gms_ -s 6 -c 4 -f -o gpurun_out/gms_full \
   python tools/ncu_gms.py > gpurun_out/ncu_gms.log 2>&1
ls -la gpurun_out | tail -20; head -12 gpurun_out/step_schnet.txt
