#!/bin/bash
# strong-scaling runs at 16384^2 under torchrun. usage: bash scripts/gpu_scale.sh TAG N [N...]
TAG=$1; shift
mkdir -p gpurun_out
export NCCL_DEBUG=WARN
for N in "$@"; do
  for RP in ${REPLS:-default}; do
    [ "$RP" = default ] && unset FSGPU_REPL_POINTS || export FSGPU_REPL_POINTS=$RP
    out=gpurun_out/${TAG}_n${N}_${RP}
    timeout ${RUN_TIMEOUT:-600} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port 29541 bench.py --gpus $N --steps 4 --warmup 3 ${BENCH_ARGS:-} > $out.json 2> $out.err
    echo "N=$N repl=$RP exit $?"
    python - <<PY
import json
try:
    d=json.loads(open("$out.json").read().strip().splitlines()[-1])
    print("   value %.1f MCell/s  ms/step %.1f  k %.2f  e2e %s  %s" % (d["value"], d["ms_per_step"], d["config"]["pcg_iters_per_solve"], d["e2e"] and round(d["e2e"]["value"],1), d["config"]["decomposition"]))
except Exception as e:
    print("   no result:", e); print(open("$out.err").read()[-1500:])
PY
  done
done
