#!/bin/bash
# Builds variants of the TMA post-trigger kernel (ring stages S, strip height TH, min CTAs/SM M, zero-row skip K) into
# build/sweep/libabub_k5t_<tag>.so; run with ABUB_LIB=<so> python tools/bench_stages.py --kernels on the GPU box.
set -e
cd "$(dirname "$0")/.."
C=autobub3hs_b200/csrc
mkdir -p build/sweep
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=false"
OBJS=$(ls $C/*.o | grep -v k5_post_tma.o)
for cfg in "$@"; do
  IFS=, read S TH M K <<< "$cfg"
  nvcc $FLAGS -DABUB_K5T_S=$S -DABUB_K5T_TH=$TH -DABUB_K5T_MINCTA=$M -DABUB_K5T_SKIP=$K -c $C/k5_post_tma.cu -o /tmp/k5t_$S$TH$M$K.o
  nvcc -shared -o build/sweep/libabub_k5t_S${S}T${TH}M${M}K${K}.so $OBJS /tmp/k5t_$S$TH$M$K.o -lcudart -ldl
  echo built S=$S TH=$TH M=$M K=$K
done
