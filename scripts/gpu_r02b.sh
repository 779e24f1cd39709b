#!/bin/bash
# Round 2, GPU call b (one GPU): whole parity suite without -x, mailbox transport on in-process ranks, smoke, bench
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_r02b.sh r02b'
TAG=${1:-r02b}; shift
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
( time timeout 900 python -m pytest tests -q -m gpu -rA --durations=15 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -3; grep -E "^(FAILED|ERROR)" gpurun_out/${TAG}_pytest.log | head -40
( time FSGPU_TEST_P2P=1 timeout 300 python -m pytest tests/test_gpu_slab.py -q -m gpu -k p2p -rA ) > gpurun_out/${TAG}_pytest_p2p.log 2>&1
echo "pytest(p2p) exit $?"; tail -15 gpurun_out/${TAG}_pytest_p2p.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${TAG}_smoke.log
( time timeout 1200 python bench.py ) > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json; tail -5 gpurun_out/${TAG}_bench.err
