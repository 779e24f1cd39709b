#!/bin/bash
# Round 2 closing call (one GPU): the driver's test command, smoke, the bench line, the other BASELINE configurations
# on the GPU and on the host, then the two ncu passes (launch list of the bench command, --set full of the level-0
# kernel classes).  usage: gpurun --timeout 1800 -- 'bash scripts/gpu_r02z.sh r02z'
TAG=${1:-r02z}
mkdir -p gpurun_out
nvidia-smi -L; nproc; free -g | head -2
( time timeout 900 python -m pytest tests -x -q -m gpu -rA --durations=10 ) > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?"; grep -E "passed|failed" gpurun_out/${TAG}_pytest.log | tail -2; grep -E "^(FAILED|ERROR)" gpurun_out/${TAG}_pytest.log | head
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"; tail -1 gpurun_out/${TAG}_smoke.log
( time timeout 1200 python bench.py ) > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"; cat gpurun_out/${TAG}_bench.json
( time timeout 600 python bench.py --impl reference --steps 5 --warmup 3 ) > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "bench(reference) exit $?"; cat gpurun_out/${TAG}_bench_ref.json
timeout 600 python scripts/bench_configs.py > gpurun_out/${TAG}_configs.jsonl 2> gpurun_out/${TAG}_configs.err; echo "configs exit $?"; cat gpurun_out/${TAG}_configs.jsonl
timeout 900 python scripts/bench_configs.py --cpu > gpurun_out/${TAG}_configs_cpu.jsonl 2> gpurun_out/${TAG}_configs_cpu.err; echo "configs(cpu) exit $?"; cat gpurun_out/${TAG}_configs_cpu.jsonl
python scripts/prof_classes.py 8192 3 > gpurun_out/${TAG}_classes.txt 2>&1; cat gpurun_out/${TAG}_classes.txt
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-ref-depth"
timeout 300 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 && timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 9000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
bash scripts/gpu_ncu2.sh ${TAG} 16
