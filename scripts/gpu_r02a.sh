#!/bin/bash
# Round 2, GPU call 1 (one GPU): the whole parity suite without -x, the mailbox transport on in-process ranks,
# the bench line, the other BASELINE configurations, the queued kernel-variant A/Bs, sanitizers on smoke().
# usage: gpurun --timeout 1500 -- 'bash scripts/gpu_r02a.sh r02a pf4 mw8 as3 as4 as6'
TAG=${1:-r02a}; shift
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
( time timeout 1500 python -m pytest tests -q -m gpu -rA --durations=15 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -3; grep -E "^(FAILED|ERROR)" gpurun_out/${TAG}_pytest.log | head -40
( time FSGPU_TEST_P2P=1 timeout 300 python -m pytest tests/test_gpu_slab.py -q -m gpu -k p2p -rA ) > gpurun_out/${TAG}_pytest_p2p.log 2>&1
echo "pytest(p2p) exit $?"; tail -8 gpurun_out/${TAG}_pytest_p2p.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -2 gpurun_out/${TAG}_smoke.log
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json
timeout 600 python scripts/bench_configs.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; echo "configs exit $?"; cat gpurun_out/${TAG}_configs.jsonl; tail -3 gpurun_out/${TAG}_configs.err
for v in "$@"; do
  FSGPU_LIB=$PWD/fluidsolver_b200/libfsgpu_$v.so timeout 600 python -m pytest tests/test_gpu_solver.py -q -m gpu -x > gpurun_out/${TAG}_parity_$v.log 2>&1
  echo "== parity $v exit $?"; tail -2 gpurun_out/${TAG}_parity_$v.log
done
bash scripts/gpu_variants.sh ${TAG} "$@"
timeout 400 compute-sanitizer --tool memcheck --error-exitcode 9 python __graft_entry__.py --smoke > gpurun_out/${TAG}_memcheck.log 2>&1; echo "memcheck exit $?"; tail -4 gpurun_out/${TAG}_memcheck.log
timeout 400 compute-sanitizer --tool racecheck --error-exitcode 9 python __graft_entry__.py --smoke > gpurun_out/${TAG}_racecheck.log 2>&1; echo "racecheck exit $?"; tail -4 gpurun_out/${TAG}_racecheck.log
