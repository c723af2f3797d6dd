mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tn_grouped.py tests/test_gpu_painn.py tests/test_gpu_dense_tc.py -x -q 2>&1 | tail -4
for w in painn schnet; do
timeout 600 python bench.py --workload $w --steps 10 --warmup 3 --no-roofline --no-cpu-baseline 2>gpurun_out/bench_$w.err | tail -1 > gpurun_out/bench_${w}_v8.json
python -c "
import json; d=json.load(open('gpurun_out/bench_${w}_v8.json')); print('$w', d['ms_per_step'], d['value'], d['e2e']['value'], d['loss'])"
done
