#!/bin/bash
# N-GPU check of the C5 exchange: NCCL test + gcn bench with both exchanges.  Usage: bash tools/gpu_halo.sh N
N=${1:-2}
mkdir -p gpurun_out
( timeout 600 python -m pytest tests/test_gpu_multi.py tests/test_gpu_gcn.py -x -q 2>&1 | tail -8 ) > gpurun_out/pytest_multi.log
for ex in halo allgather; do
  ( timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 10 --warmup 3 --workload gcn --gcn-exchange $ex --no-cpu-baseline 2>gpurun_out/gcn_${ex}_$N.err | tail -1 ) > gpurun_out/gcn_${ex}_$N.json
done
( timeout 300 python bench.py --steps 10 --warmup 3 --workload gcn --no-cpu-baseline 2>gpurun_out/gcn_1.err | tail -1 ) > gpurun_out/gcn_1.json
cat gpurun_out/pytest_multi.log; for f in gpurun_out/gcn_halo_$N gpurun_out/gcn_allgather_$N gpurun_out/gcn_1; do python - $f.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); print(sys.argv[1], "ms/step %.3f" % d["ms_per_step"], "e2e ms %.3f" % d["e2e"]["ms_per_step"], d["config"].get("exchange"), d["config"]["parallelism"][:150])
except Exception as e:
    print(sys.argv[1], "FAILED", e)
PY
tail -3 $f.err; done
