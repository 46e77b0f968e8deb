#!/bin/bash
# Builds variants of the TMA trigger kernel (rows per box R, ring stages S, min CTAs/SM) into
# build/sweep/libabub_<tag>.so; run on the GPU box with tools/k1_sweep_run.sh.
set -e
cd "$(dirname "$0")/.."
C=autobub3hs_b200/csrc
mkdir -p build/sweep
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC --fmad=false"
OBJS="$C/ctx.o $C/train.o $C/k1_processframe.o $C/k2_trigger_scan.o $C/k34_trainer.o $C/k5_localize.o"
for cfg in "$@"; do
  IFS=, read R S M <<< "$cfg"
  nvcc $FLAGS -DABUB_K1T_R=$R -DABUB_K1T_S=$S -DABUB_K1T_MINCTA=$M -c $C/k1_trigger_tma.cu -o /tmp/k1t_$R$S$M.o
  nvcc -shared -o build/sweep/libabub_R${R}S${S}M${M}.so $OBJS /tmp/k1t_$R$S$M.o -lcudart -ldl
  echo built R=$R S=$S M=$M
done
