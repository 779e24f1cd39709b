#!/bin/bash
# parity tests + smoke only (cheap GPU call). usage: bash scripts/gpu_tests.sh [tag] [pytest args]
TAG=${1:-t}; shift
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -q -rA --durations=15 "$@" > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -40 gpurun_out/${TAG}_pytest.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
tail -5 gpurun_out/${TAG}_smoke.log
