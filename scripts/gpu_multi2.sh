#!/bin/bash
# multi-GPU comparison of the transports at 16384^2. usage: bash scripts/gpu_multi2.sh TAG N "variant variant ..."
# variants: p2p (mailbox, full bench line), p2p_lite (no anchor / e2e), p2p_ovl (boundary-first overlap), nccl, nccl_full
TAG=$1; N=$2; shift 2
mkdir -p gpurun_out
nvidia-smi -L | head -8
export NCCL_DEBUG=WARN
port=29541
for VV in ${@:-p2p p2p_noovl nccl}; do
  # a variant may carry environment settings: name:ENV=VAL,ENV=VAL
  V=${VV%%:*}; ENVS=""; [ "$VV" != "$V" ] && ENVS=${VV#*:}
  unset FSGPU_OVERLAP FSGPU_REPL_POINTS FSGPU_OVERLAP_MIN_POINTS
  for kv in ${ENVS//,/ }; do export "$kv"; done
  case $V in
    p2p) ARGS="--halo p2p" ;;
    p2p_ovl) ARGS="--halo p2p --no-anchor --no-e2e --no-ref-depth"; export FSGPU_OVERLAP=1 ;;
    p2p_lite) ARGS="--halo p2p --no-anchor --no-e2e --no-ref-depth" ;;
    nccl_full) ARGS="--halo nccl" ;;
    nccl) ARGS="--halo nccl --no-anchor --no-e2e --no-ref-depth" ;;
    *) ARGS="$V" ;;
  esac
  out=gpurun_out/${TAG}_n${N}_${VV//[:=,]/_}
  port=$((port+1))
  ( time timeout ${RUN_TIMEOUT:-240} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 \
      --master-port $port bench.py --gpus $N --steps ${STEPS:-5} --warmup 3 $ARGS ) > $out.json 2> $out.err
  echo "== N=$N $VV exit $?"
  python - <<PY
import json
try:
    d=json.loads([l for l in open("$out.json").read().strip().splitlines() if l.startswith("{")][-1])
    a=d["config"].get("same_problem_1gpu")
    print("   value %.1f MCell/s  ms/step %.1f  k %.2f  e2e %s  maxU %.6g\n   %s" % (d["value"], d["ms_per_step"], d["config"]["pcg_iters_per_solve"], d["e2e"] and round(d["e2e"]["value"],1), d["config"]["max_abs_U"], d["config"]["decomposition"]))
    if a: print("   1-GPU anchor:", a, " -> efficiency %.3f" % (d["value"]/a["value"]/d["n_gpus"]) if "value" in a else "")
except Exception as e:
    print("   no result:", e); print(open("$out.err").read()[-2500:])
PY
  grep -i "fsgpu\]\|error\|timed out" $out.err | head -5
done
