#!/bin/bash
# Profiling recipe run on the GPU box through gpurun (B200_PROFILING.md): launch list of the bench command, one
# `ncu --set full` capture of one view's kernels, and one of the auxiliary kernels (k-buffer pair, model-side passes).
# Usage: scripts/profile.sh <tag>     then here: scripts/ncu_summarise.py <tag>
set -u
TAG=${1:-r01}
OUT=gpurun_out
mkdir -p $OUT
BENCH="python bench.py --steps 2 --warmup 3 --no-e2e --no-cpu-baseline"
ONE="python bench.py --steps 1 --warmup 3 --views 1 --no-e2e --no-cpu-baseline"
$BENCH > $OUT/plain_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_${TAG}.csv $BENCH > $OUT/ncu_launches_${TAG}.log 2>&1
echo "launch list rc=$?"
# the same without the auxiliary measurements (forward sweeps, k-buffer, model-side passes, whole training step): the launches of
# the timed steps and their warm-up only, so that a kernel's share here is its share of a step
STEP="$BENCH --no-aux --no-sweep"
$STEP > $OUT/plain_step_${TAG}.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $OUT/launches_${TAG}_step.csv $STEP > $OUT/ncu_launches_step_${TAG}.log 2>&1
echo "step launch list rc=$?"
$ONE --no-aux > $OUT/plain_one_${TAG}.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:gut_ -s 33 -c 11 -f -o $OUT/prof_${TAG} $ONE --no-aux > $OUT/ncu_full_${TAG}.log 2>&1
echo "full capture rc=$?"
# (gpurun brings back at most 64 MiB: the auxiliary captures carry no source and, for the model-side passes, only the summarised metrics)
ncu --set full --clock-control none -k regex:gut_render_kbuffer -s 2 -c 2 -f -o $OUT/prof_${TAG}_kbuffer $ONE > $OUT/ncu_kbuffer_${TAG}.log 2>&1
echo "k-buffer capture rc=$?"
M=gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed,sm__throughput.avg.pct_of_peak_sustained_elapsed,smsp__inst_executed.sum,launch__registers_per_thread,sm__warps_active.avg.pct_of_peak_sustained_active,launch__grid_size,launch__block_size
ncu --metrics $M --clock-control none -k "regex:gut_selective_adam|gut_activate|gut_pack_features|gut_unpack_features|gut_densify" -c 36 -f -o $OUT/prof_${TAG}_model $ONE > $OUT/ncu_model_${TAG}.log 2>&1
echo "model-side capture rc=$?"
ls -la $OUT | grep ${TAG}; du -sh $OUT
