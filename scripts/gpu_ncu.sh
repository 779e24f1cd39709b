#!/bin/bash
# ncu passes on the bench command (after it has exited 0 without ncu). usage: bash scripts/gpu_ncu.sh TAG [skip] [count]
TAG=${1:-ncu}; SKIP=${2:-1500}; CNT=${3:-45}
mkdir -p gpurun_out
BCMD="python bench.py --steps 1 --warmup 1 --no-cpu --no-e2e --no-ref-depth"
timeout 300 $BCMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 8000 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $BCMD > gpurun_out/${TAG}_ncu1.log 2>&1
echo "ncu launches exit $?"
timeout 900 ncu --set full --clock-control none --import-source on -s $SKIP -c $CNT \
    -o /tmp/${TAG}_full -f $BCMD > gpurun_out/${TAG}_ncu2.log 2>&1
echo "ncu full exit $?"
if [ -f /tmp/${TAG}_full.ncu-rep ]; then
  ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
  SZ=$(stat -c %s /tmp/${TAG}_full.ncu-rep); echo "rep size $SZ"
  [ "$SZ" -lt 45000000 ] && cp /tmp/${TAG}_full.ncu-rep gpurun_out/
fi
ls -la gpurun_out | tail -8
