#!/bin/bash
# Build a variant of libgut_b200 for A/B timing: scripts/ab_build.sh <tag> [extra nvcc flags for gut_render.cu, e.g. -DGUT_BWD_MIN_BLOCKS=24]
# -> gpurun_out/ is NOT shipped to the box, so variants go to build_ab/libgut_<tag>.so (git-ignored via build*/ and *.so).
set -eu
TAG=$1; shift
cd "$(dirname "$0")/../3dgrut_b200/csrc"
mkdir -p ../../build_ab
ARCH="-gencode arch=compute_100a,code=sm_100a"
nvcc $ARCH -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -Xptxas -v -use_fast_math "$@" -c gut_render.cu -o ../../build_ab/gut_render_$TAG.o 2> ../../build_ab/gut_render_$TAG.ptxas.log
nvcc $ARCH -shared -o ../../build_ab/libgut_$TAG.so gut_api.o gut_project.o gut_binning.o ../../build_ab/gut_render_$TAG.o gut_render_kbuffer.o gut_model.o -lcudart
grep -A3 "gut_render_backward_kernelILi2ELb1" ../../build_ab/gut_render_$TAG.ptxas.log | grep -i "spill\|Used" 
