#!/bin/bash
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_cfconv.py tests/test_gpu_schnet.py tests/test_gpu_capacity.py tests/test_gpu_parity_full.py -x -q 2>&1 | tail -3 ) > gpurun_out/pytest_q.log
tail -2 gpurun_out/pytest_q.log
( timeout 300 python bench.py --workload schnet --steps 20 --warmup 5 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_q.err | tail -1 ) > gpurun_out/bench_q.json
python -c "
import json
d=json.load(open('gpurun_out/bench_q.json')); print('schnet ms/step %.3f' % d['ms_per_step'], 'value %.0f' % d['value'], 'e2e %.0f' % d['e2e']['value'])" || tail -3 gpurun_out/bench_q.err
timeout 300 python tools/profile_step.py schnet --batch 256 --rows 16 > gpurun_out/step_q.txt 2>/dev/null; head -18 gpurun_out/step_q.txt | cut -c1-130
