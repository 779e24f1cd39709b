#!/bin/bash
# tests + per-class timing with and without the one-kernel multigrid tail. usage: bash scripts/gpu_ab2.sh TAG
TAG=${1:-ab}
mkdir -p gpurun_out
( timeout 900 python -m pytest tests -x -q -m gpu ) > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest exit $?"; tail -3 gpurun_out/${TAG}_pytest.log
python scripts/prof_classes.py 8192 3 > gpurun_out/${TAG}_classes.txt 2>&1; cat gpurun_out/${TAG}_classes.txt
FSGPU_NO_TAIL=1 python scripts/prof_classes.py 8192 3 > gpurun_out/${TAG}_classes_notail.txt 2>&1; cat gpurun_out/${TAG}_classes_notail.txt
