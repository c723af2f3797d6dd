#!/bin/bash
# full GPU parity suite + smoke
mkdir -p gpurun_out
( timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -25 ) > gpurun_out/pytest_gpu.log
( timeout 200 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -5 ) > gpurun_out/smoke.log
tail -25 gpurun_out/pytest_gpu.log; cat gpurun_out/smoke.log
