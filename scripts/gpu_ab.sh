#!/bin/bash
# A/B of library builds: parity tests on the default build, then short benches per build and
# kernel class.  usage: bash scripts/gpu_ab.sh TAG lib1.so lib2.so ...
TAG=$1; shift
mkdir -p gpurun_out
if [ -z "${NOTEST:-}" ]; then
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_pytest.log
fi
for lib in "$@"; do
  for prof in ${PROFS:-vcycle_up_l0 vcycle_down_l0 matvec_dot xr_update}; do
    FSGPU_LIB=$PWD/fluidsolver_b200/$lib timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-e2e --prof $prof \
      > gpurun_out/${TAG}_${lib%.so}_${prof}.json 2>> gpurun_out/${TAG}_err.log
    python - <<PY
import json
d=json.load(open("gpurun_out/${TAG}_${lib%.so}_${prof}.json"))
r=d["roofline"]
print("${lib}", "${prof}", "value %.1f ms/step %.1f k %.2f | kernel avg %.1f us frac %.3f share %.3f step_frac %.3f" % (d["value"], d["ms_per_step"], d["config"]["pcg_iters_per_solve"], r["avg_launch_us"], r["frac"], r["share_of_step"], r["step_frac"]))
PY
  done
done
