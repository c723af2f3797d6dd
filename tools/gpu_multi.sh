#!/bin/bash
# N-GPU bench pass (launched the way the driver does it). Usage: bash tools/gpu_multi.sh N
N=${1:-2}
mkdir -p gpurun_out
for w in schnet hdnnp gcn painn; do
  ( timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus $N --steps 10 --warmup 3 --workload $w --no-cpu-baseline 2>gpurun_out/multi_${w}_$N.err | tail -1 ) > gpurun_out/multi_${w}_$N.json
done
( timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 \
      bench.py --impl reference --gpus $N --steps 2 --warmup 1 2>&1 | tail -1 ) > gpurun_out/multi_ref_$N.json
cat gpurun_out/multi_*_$N.json | cut -c1-700; for f in gpurun_out/multi_*_$N.err; do tail -2 $f; done
