#!/bin/bash
# Builds gpt_b200/lib/libgptb200.so for sm_100a (cross-compiles without a GPU).
# GPTB_OUT=<dir> EXTRA_NVCC_FLAGS="-D..." builds an experiment variant into <dir> (select it with GPTB200_LIB).
set -e
cd "$(dirname "$0")"
OUT=${GPTB_OUT:-gpt_b200/lib}
mkdir -p "$OUT"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -ccbin /usr/bin/g++ $EXTRA_NVCC_FLAGS"
OBJS=""
pids=""
for f in api batch classify comm contract contract_tc fp64_path fulltheta gmc multi nodes pred_nodes tc_path tgp update; do
  $NVCC $FLAGS -c gpt_b200/csrc/$f.cu -o "$OUT/$f.o" &
  pids="$pids $!"
  OBJS="$OBJS $OUT/$f.o"
done
for p in $pids; do wait $p; done
$NVCC -shared -o "$OUT/libgptb200.so" $OBJS -gencode arch=compute_100a,code=sm_100a -ccbin /usr/bin/g++ -ldl
echo built "$OUT/libgptb200.so"
