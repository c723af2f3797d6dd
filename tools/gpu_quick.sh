#!/bin/bash
# quick GPU pass: selected tests (args = pytest -k expression) + smoke
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -m gpu -x -q -k "$1" 2>&1 | tail -30 ) > gpurun_out/pytest_quick.log
( timeout 300 python -c 'import __graft_entry__ as g; g.smoke()' 2>&1 | tail -5 ) > gpurun_out/smoke.log
tail -30 gpurun_out/pytest_quick.log; cat gpurun_out/smoke.log
