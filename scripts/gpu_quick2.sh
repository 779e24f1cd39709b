#!/bin/bash
# whole GPU suite + per-class timings at 8192^2. usage: bash scripts/gpu_quick2.sh TAG
TAG=${1:-q}
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests -q -m gpu -rA --durations=8 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -2; grep -E "^(FAILED|ERROR)" gpurun_out/${TAG}_pytest.log | head -20
timeout 300 python scripts/prof_classes.py 8192 3 > gpurun_out/${TAG}_prof.txt 2>&1; cat gpurun_out/${TAG}_prof.txt
