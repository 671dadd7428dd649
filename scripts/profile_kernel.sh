#!/bin/bash
# ncu --set full capture of ONE kernel (regex $1) of a one-view bench run; output gpurun_out/prof_$2.ncu-rep
set -u
K=$1; TAG=$2; SKIP=${3:-3}
ONE="python bench.py --steps 1 --warmup 3 --views 1 --no-e2e --no-cpu-baseline"
$ONE > gpurun_out/plain_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c 1 -f -o gpurun_out/prof_${TAG} $ONE > gpurun_out/ncu_${TAG}.log 2>&1
echo "rc=$?"
