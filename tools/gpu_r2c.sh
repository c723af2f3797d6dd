#!/bin/bash
# full GPU test suite + dense microbench + schnet/painn bench
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -15 ) > gpurun_out/pytest_gpu.log
( timeout 600 python tools/mb_dense.py 2>&1 | tail -20 ) > gpurun_out/mb_dense.txt
( timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_schnet.err | tail -2 ) > gpurun_out/bench_schnet.json
( timeout 600 python bench.py --steps 5 --warmup 3 --workload painn --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_painn.err | tail -2 ) > gpurun_out/bench_painn.json
( timeout 600 python bench.py --steps 5 --warmup 3 --workload hdnnp --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/bench_hdnnp.err | tail -2 ) > gpurun_out/bench_hdnnp.json
tail -15 gpurun_out/pytest_gpu.log; cat gpurun_out/mb_dense.txt; for w in schnet painn hdnnp; do cut -c1-250 gpurun_out/bench_$w.json; done
