#!/bin/bash
# how often do the in-process mailbox tests hit a device-side timeout?  usage: bash scripts/gpu_flaky.sh N
N=${1:-8}
mkdir -p gpurun_out
for ML in LAZY EAGER; do
  fails=0; retries=0
  for i in $(seq 1 $N); do
    CUDA_MODULE_LOADING=$ML FSGPU_TEST_OVERLAP=1 timeout 200 python -m pytest tests/test_gpu_slab.py tests/test_gpu_twophase.py tests/test_gpu_zzz_inprocess_mailbox.py -q -m gpu -s > gpurun_out/flaky_${ML}_$i.log 2>&1 || fails=$((fails+1))
    r=$(grep -c "retrying once" gpurun_out/flaky_${ML}_$i.log); retries=$((retries+r))
  done
  echo "CUDA_MODULE_LOADING=$ML: $N runs, $fails failed, $retries retries"
done
