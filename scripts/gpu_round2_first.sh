#!/bin/bash
# First GPU call of the next session: everything that was written after round 1's GPU budget ran out,
# in the order that decides most per minute.  Build the variant libraries HERE first (nvcc cross-compiles):
#   python scripts/build_variants.py pf4=solver_cycle.cu:-DFS_PF_ROWS=4 mw8=solver_cycle.cu:-DFS_MWARPS=8 \
#       as3=solver_cycle.cu:-DFS_UP_ASYNC=3,-DFS_DOWN_ASYNC=3 as4=solver_cycle.cu:-DFS_UP_ASYNC=4,-DFS_DOWN_ASYNC=4 \
#       as6=solver_cycle.cu:-DFS_UP_ASYNC=6,-DFS_DOWN_ASYNC=6,-DFS_MARCH_MINB=4
# then:  gpurun --timeout 1500 -- 'bash scripts/gpu_round2_first.sh r02a pf4 mw8 as3 as4 as6'
TAG=${1:-r02a}; shift
mkdir -p gpurun_out
# 1. the last-sorted test file first (all post-budget rows), without -x: every failure is wanted at once
( time timeout 600 python -m pytest tests/test_gpu_zz_driver_variants.py -q -m gpu --durations=10 ) > gpurun_out/${TAG}_pytest_zz.log 2>&1
echo "pytest(zz) exit $?"; tail -25 gpurun_out/${TAG}_pytest_zz.log
# 2. the whole suite as the driver runs it
( time timeout 1500 python -m pytest tests -x -q -m gpu --durations=10 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; tail -8 gpurun_out/${TAG}_pytest.log
# 3. mailbox transport, in-process ranks on one GPU
( time FSGPU_TEST_P2P=1 timeout 300 python -m pytest tests/test_gpu_slab.py -q -m gpu -k p2p ) > gpurun_out/${TAG}_pytest_p2p.log 2>&1
echo "pytest(p2p) exit $?"; tail -8 gpurun_out/${TAG}_pytest_p2p.log
# 4. bench line, then the other configurations
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json
timeout 600 python scripts/bench_configs.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; echo "configs exit $?"; cat gpurun_out/${TAG}_configs.jsonl
# 5. kernel variants: parity of each library first (solver tests under FSGPU_LIB), then per-class timing
for v in "$@"; do
  FSGPU_LIB=$PWD/fluidsolver_b200/libfsgpu_$v.so timeout 600 python -m pytest tests/test_gpu_solver.py -q -m gpu -x > gpurun_out/${TAG}_parity_$v.log 2>&1
  echo "== parity $v exit $?"; tail -2 gpurun_out/${TAG}_parity_$v.log
done
bash scripts/gpu_variants.sh ${TAG} "$@"
# 6. the convergence study incl. the three sizes the reference leaves commented out
timeout 900 python examples/channel_convergence.py --csv --max-n 9 > gpurun_out/${TAG}_study.csv 2> gpurun_out/${TAG}_study.err; echo "study exit $?"; cat gpurun_out/${TAG}_study.csv
