mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attention.py tests/test_gpu_graph.py -x -q 2>&1 | tail -25
timeout 600 python bench.py --workload hdnnp --steps 10 --warmup 3 --no-roofline 2>gpurun_out/bench_hdnnp.err | tail -1 > gpurun_out/bench_hdnnp_devangles.json
cat gpurun_out/bench_hdnnp_devangles.json; tail -5 gpurun_out/bench_hdnnp.err
