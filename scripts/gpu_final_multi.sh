#!/bin/bash
# closing multi-GPU call: parity of the slab path against the 1-GPU run on N real GPUs, then the full bench line
# (library defaults) as the driver launches it.  usage: gpurun --gpus N -- 'bash scripts/gpu_final_multi.sh TAG N'
TAG=$1; N=$2
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29651 \
  scripts/multi_parity.py ${PGRID:-4096} 2 > gpurun_out/${TAG}_parity_n$N.txt 2> gpurun_out/${TAG}_parity_n$N.err
echo "parity exit $?"; cat gpurun_out/${TAG}_parity_n$N.txt; grep -i "error\|timed out" gpurun_out/${TAG}_parity_n$N.err | head -3
( time timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29652 \
  bench.py --gpus $N --steps 5 --warmup 3 ) > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
echo "bench exit $?"; cat gpurun_out/${TAG}_bench_n$N.json
