#!/bin/bash
# multi-GPU bench under torchrun. usage: bash scripts/gpu_multi.sh TAG N [extra bench args]
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
run() { # name, nproc, args...
  local name=$1 np=$2; shift 2
  if [ "$np" = 1 ]; then
    timeout 600 python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 \
      --master-port 29533 bench.py --gpus $np "$@" > gpurun_out/${TAG}_${name}.json 2> gpurun_out/${TAG}_${name}.err
  fi
  echo "$name exit $?"; tail -c 1500 gpurun_out/${TAG}_${name}.json; tail -5 gpurun_out/${TAG}_${name}.err
}
nvidia-smi -L
run n1_4096 1 --grid 4096 --steps 2 --warmup 1 --no-cpu --no-e2e
run n${N}_4096 $N --grid 4096 --steps 2 --warmup 1 --no-cpu --no-e2e
run n${N}_full $N --steps 3 --warmup 2 "$@"
