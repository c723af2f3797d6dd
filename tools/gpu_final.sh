#!/bin/bash
# Round-end pass: full GPU parity suite, smoke, step tables (PDL off so that CUPTI durations do not include the
# dependency wait), default bench line + reference arm, ncu launch list of the default bench command.
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -6 ) > gpurun_out/pytest_gpu.log
( timeout 200 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -3 ) > gpurun_out/smoke.log
for w in schnet painn hdnnp; do
  b=256; [ $w = painn ] && b=512
  KGCNN_PDL=0 timeout 600 python tools/profile_step.py $w --batch $b --rows 70 > gpurun_out/step_$w.txt 2>gpurun_out/step_$w.err
done
( timeout 900 python bench.py 2>gpurun_out/bench_default.err | tail -1 ) > gpurun_out/bench_default.json
( timeout 600 python bench.py --impl reference --steps 3 --warmup 1 2>gpurun_out/bench_ref.err | tail -1 ) > gpurun_out/bench_ref.json
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_default.csv \
   python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads > gpurun_out/ncu_launch.log 2>&1
tail -4 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log | cut -c1-200; head -3 gpurun_out/step_*.txt | cut -c1-120; cut -c1-400 gpurun_out/bench_default.json; cut -c1-300 gpurun_out/bench_ref.json; wc -l gpurun_out/launches_default.csv
