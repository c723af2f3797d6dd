mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hdnnp.py tests/test_gpu_wacsf.py -x -q 2>&1 | tail -15
timeout 600 python bench.py --workload hdnnp --steps 10 --warmup 3 --no-roofline 2>gpurun_out/bench_hdnnp.err | tail -1 > gpurun_out/bench_hdnnp_v5.json
python -c "
import json; d=json.load(open('gpurun_out/bench_hdnnp_v5.json')); print(d['ms_per_step'], d['value'], d['e2e'], d['loss'])"
tail -3 gpurun_out/bench_hdnnp.err
timeout 600 python tools/profile_step.py hdnnp --batch 256 --rows 14 2>/dev/null | tee gpurun_out/step_hdnnp_v5.txt | head -16
