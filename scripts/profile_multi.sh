#!/bin/bash
# ncu --set full of the LAST launch of several kernels in a 2-view run (second view = accumulate mode)
set -u
TAG=$1; shift
CMD="python bench.py --steps 1 --warmup 1 --views 2 --no-e2e --no-cpu-baseline"
$CMD > gpurun_out/plain_${TAG}.log 2>&1 || { echo plain failed; exit 1; }
for K in "$@"; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 3 -c 1 -f -o gpurun_out/prof_${TAG}_$K $CMD > gpurun_out/ncu_${TAG}_$K.log 2>&1
  echo "$K rc=$?"
done
