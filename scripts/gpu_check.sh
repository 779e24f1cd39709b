#!/bin/bash
# One GPU-box pass: parity tests, bench line, ncu launch list, ncu --set full of the top kernels.
# usage (under gpurun): bash scripts/gpu_check.sh [tag]
set -u
TAG=${1:-r01}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/${TAG}_smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1
echo "pytest exit $?" >> gpurun_out/${TAG}_pytest.log
tail -3 gpurun_out/${TAG}_pytest.log
timeout 600 python __graft_entry__.py --smoke > gpurun_out/${TAG}_smoke.log 2>&1; echo "smoke exit $?"
timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err; echo "bench exit $?"
cat gpurun_out/${TAG}_bench.json
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e"
timeout 300 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
# full capture of one PCG iteration's kernels; the .ncu-rep stays on the box if it is large
# (gpurun_out is capped at 64 MiB), its raw page always comes back as CSV
timeout 300 $BCMD > gpurun_out/${TAG}_plain2.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -s 3000 -c 40 \
    -o /tmp/${TAG}_full -f $BCMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
if [ -f /tmp/${TAG}_full.ncu-rep ]; then
  ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
  SZ=$(stat -c %s /tmp/${TAG}_full.ncu-rep)
  [ "$SZ" -lt 45000000 ] && cp /tmp/${TAG}_full.ncu-rep gpurun_out/
fi
ls -la gpurun_out
# strong-scaling baseline: the 16384^2 workload on ONE GPU
timeout 600 python bench.py --grid 16384 --steps 3 --warmup 2 --no-cpu > gpurun_out/${TAG}_bench16384.json 2> gpurun_out/${TAG}_bench16384.err; echo "bench16384 exit $?"
cat gpurun_out/${TAG}_bench16384.json
