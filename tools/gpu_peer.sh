#!/bin/bash
# N-GPU check of the peer-memory all-reduce: test, then the C2 bench with it on / off.  Usage: bash tools/gpu_peer.sh N
N=${1:-2}
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_multi.py -x -q 2>&1 | tail -12 ) > gpurun_out/pytest_peer.log; cat gpurun_out/pytest_peer.log
for peer in auto 1 0; do
  ( KGCNN_PEER_ALLREDUCE=$peer timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 \
      bench.py --gpus $N --steps 30 --warmup 5 --workload schnet --no-cpu-baseline --no-roofline --no-other-workloads 2>gpurun_out/peer_${peer}_$N.err | tail -1 ) > gpurun_out/peer_${peer}_$N.json
  python -c "
import json
d=json.load(open('gpurun_out/peer_${peer}_$N.json')); print('PEER=$peer N=$N ms/step %.3f' % d['ms_per_step'], 'value %.0f' % d['value'], 'loss', d.get('loss'))" || grep -v "OMP_NUM\|\*\*\*" gpurun_out/peer_${peer}_$N.err | tail -15
done
