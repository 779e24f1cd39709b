#!/bin/bash
# TMA ring kernels: bit-identity test, then per-class timings ring on / off at 8192^2
TAG=${1:-ring}
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_solver.py -q -m gpu -k "ring" -x -rA ) > gpurun_out/${TAG}_test.log 2>&1
echo "ring test exit $?"; tail -15 gpurun_out/${TAG}_test.log
[ "$2" = "testonly" ] && exit 0
for R in 1 0; do
  FSGPU_RING=$R timeout 300 python scripts/prof_classes.py 8192 2 > gpurun_out/${TAG}_prof_ring$R.txt 2>&1
  echo "== FSGPU_RING=$R"; cat gpurun_out/${TAG}_prof_ring$R.txt | tail -16
done
