#!/bin/bash
# per-operation cost of the slab transports. usage: bash scripts/gpu_probe.sh TAG N [grid]
TAG=$1; N=$2; GRID=${3:-16384}
mkdir -p gpurun_out
port=29700
for H in nccl p2p; do
  port=$((port+1))
  FSGPU_HALO=$H timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
    --master-port $port scripts/comm_probe.py $GRID > gpurun_out/${TAG}_probe_n${N}_$H.txt 2> gpurun_out/${TAG}_probe_n${N}_$H.err
  echo "== N=$N $H exit $?"; cat gpurun_out/${TAG}_probe_n${N}_$H.txt
done
