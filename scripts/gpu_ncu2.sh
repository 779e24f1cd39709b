#!/bin/bash
# ncu --set full on the level-0 kernels only (classes momentum, rhs, correct, matvec, xr, pupd, down0, up0 = mask 0x3bd)
# usage: bash scripts/gpu_ncu2.sh TAG [count]
TAG=${1:-ncu2}; CNT=${2:-14}
mkdir -p gpurun_out
CMD="python scripts/ncu_step.py 8192"
timeout 300 $CMD > gpurun_out/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/${TAG}_plain.log; exit 1; }
FSGPU_NCU_CLASSES=0x3bd timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -c $CNT \
    -o /tmp/${TAG}_full -f $CMD > gpurun_out/${TAG}_ncu.log 2>&1
echo "ncu full exit $?"; tail -3 gpurun_out/${TAG}_ncu.log
if [ -f /tmp/${TAG}_full.ncu-rep ]; then
  ncu -i /tmp/${TAG}_full.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_raw.csv 2>/dev/null
  SZ=$(stat -c %s /tmp/${TAG}_full.ncu-rep); echo "rep size $SZ"
  [ "$SZ" -lt 60000000 ] && cp /tmp/${TAG}_full.ncu-rep gpurun_out/
fi
