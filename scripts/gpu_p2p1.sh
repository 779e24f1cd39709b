#!/bin/bash
# mailbox transport on in-process ranks (one GPU): the p2p tests, then the whole slab file
TAG=${1:-p2p1}
mkdir -p gpurun_out
( time timeout 300 python -m pytest tests/test_gpu_slab.py -q -m gpu -k p2p -rA -x ) > gpurun_out/${TAG}_p2p.log 2>&1
echo "pytest(p2p) exit $?"; grep -E "passed|failed|Error|timed out" gpurun_out/${TAG}_p2p.log | tail -12
( time timeout 300 python -m pytest tests/test_gpu_slab.py tests/test_gpu_solver.py -q -m gpu ) > gpurun_out/${TAG}_slab.log 2>&1
echo "pytest(slab+solver) exit $?"; tail -3 gpurun_out/${TAG}_slab.log
