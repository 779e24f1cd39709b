#!/bin/bash
# effect of the multigrid depth (the reference caps PFMG at 16 levels) on iterations and step time
TAG=$1; shift
mkdir -p gpurun_out
for ML in "$@"; do
  timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --max-levels $ML ${BENCH_ARGS:-} > gpurun_out/${TAG}_ml${ML}.json 2>> gpurun_out/${TAG}_err.log
  python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_ml${ML}.json"))
print("max_levels ${ML}: levels %s value %.1f ms/step %.1f k %.2f step_frac %.3f" % (d["config"]["mg_levels"], d["value"], d["ms_per_step"], d["config"]["pcg_iters_per_solve"], d["roofline"]["step_frac"]))
PY
done
