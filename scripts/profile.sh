#!/bin/bash
# Profiling recipe run on the GPU box through gpurun (B200_PROFILING.md): launch list of the bench command, then one
# `ncu --set full` capture of one view's kernels.  Usage: scripts/profile.sh <tag>
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
ONE="python bench.py --steps 1 --warmup 3 --views 1 --no-e2e --no-cpu-baseline"
$BENCH > $OUT/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_${TAG}.csv $BENCH > $OUT/ncu_launches_${TAG}.log 2>&1
echo "launch list rc=$?"
$ONE > $OUT/plain_one_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gut_ -s 33 -c 11 -f -o $OUT/prof_${TAG} $ONE > $OUT/ncu_full_${TAG}.log 2>&1
echo "full capture rc=$?"
ls -la $OUT
