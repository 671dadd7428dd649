#!/bin/bash
# A/B timing of library variants built by scripts/ab_build.sh: scripts/ab_run.sh <tag> [<tag> ...]   ("base" = the in-tree library)
mkdir -p gpurun_out
for t in "$@"; do
  if [ "$t" = base ]; then unset GUT_B200_LIBRARY; else export GUT_B200_LIBRARY=$PWD/build_ab/libgut_$t.so; fi
  echo -n "$t: "; scripts/quick_bench.sh $t
done
