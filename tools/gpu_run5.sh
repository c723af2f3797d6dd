mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hdnnp.py tests/test_gpu_wacsf.py -x -q 2>&1 | tail -8
timeout 600 python bench.py --workload hdnnp --steps 10 --warmup 3 --no-roofline 2>gpurun_out/bench_hdnnp.err | tail -1 > gpurun_out/bench_hdnnp_v7.json
python -c "
import json; d=json.load(open('gpurun_out/bench_hdnnp_v7.json')); print(d['ms_per_step'], d['value'], d['e2e'], d['loss'])"
timeout 600 python tools/ncu_g4.py > gpurun_out/g4_plain.log 2>&1; tail -1 gpurun_out/g4_plain.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:acsf_g4 -c 3 -f -o gpurun_out/g4_full python tools/ncu_g4.py > gpurun_out/ncu_g4.log 2>&1; tail -2 gpurun_out/ncu_g4.log
