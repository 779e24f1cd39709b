#!/bin/bash
# per-class device time under slab decomposition. usage: bash scripts/gpu_prof_multi.sh TAG N [halo ...]
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
port=29600
for H in ${@:-p2p nccl}; do
  port=$((port+1))
  FSGPU_HALO=$H timeout ${RUN_TIMEOUT:-200} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $port scripts/prof_classes.py 16384 2 > gpurun_out/${TAG}_prof_n${N}_$H.txt 2> gpurun_out/${TAG}_prof_n${N}_$H.err
  echo "== N=$N halo=$H exit $?"; cat gpurun_out/${TAG}_prof_n${N}_$H.txt
done
