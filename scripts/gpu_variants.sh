#!/bin/bash
# per-class timing of the default library and of each variant library given. usage: bash scripts/gpu_variants.sh TAG v1 v2 ...
TAG=$1; shift
mkdir -p gpurun_out
python scripts/prof_classes.py 8192 2 > gpurun_out/${TAG}_base.txt 2>&1; echo "== base"; grep -v "bconds\|coarsest\|div_rhs\|shift" gpurun_out/${TAG}_base.txt
for v in "$@"; do
  FSGPU_LIB=$PWD/fluidsolver_b200/libfsgpu_$v.so python scripts/prof_classes.py 8192 2 > gpurun_out/${TAG}_$v.txt 2>&1
  echo "== $v"; grep "step\|vcycle" gpurun_out/${TAG}_$v.txt
done
