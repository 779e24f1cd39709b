#!/bin/bash
# GPU-box pass without profilers: parity tests (new files first), smoke, one bench line.
# usage (under gpurun): bash scripts/gpu_quick.sh [tag]
TAG=${1:-q}
mkdir -p gpurun_out
( time timeout 900 python -m pytest tests/test_gpu_twophase.py tests/test_gpu_shim.py -q -m gpu ) > gpurun_out/${TAG}_pytest_new.log 2>&1
echo "pytest(new) exit $?" | tee -a gpurun_out/${TAG}_pytest_new.log
tail -30 gpurun_out/${TAG}_pytest_new.log
( time timeout 1500 python -m pytest tests -x -q -m gpu --durations=15 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" | tee -a gpurun_out/${TAG}_pytest.log
tail -30 gpurun_out/${TAG}_pytest.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
tail -3 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
