mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_primitives.py tests/test_gpu_schnet.py tests/test_gpu_gcn.py -x -q 2>&1 | tail -4
for w in schnet gcn; do
timeout 600 python bench.py --workload $w --steps 20 --warmup 5 --no-roofline --no-cpu-baseline 2>gpurun_out/bench_$w.err | tail -1 > gpurun_out/bench_${w}_v10.json
python -c "
import json; d=json.load(open('gpurun_out/bench_${w}_v10.json')); print('$w', d['ms_per_step'], d['value'], d['e2e']['value'], d['loss'])"
done
timeout 600 python tools/profile_step.py schnet --batch 256 --rows 12 2>/dev/null | tee gpurun_out/step_schnet_v10.txt | head -12 | cut -c1-130
