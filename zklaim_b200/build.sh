#!/bin/bash
# Builds zklaim_b200/libzkb.so (CUDA kernels + C ABI) for sm_100a.  nvcc cross-compiles without a GPU.
set -e
cd "$(dirname "$0")/csrc"
OUT=../libzkb.so
OBJ=../_obj
mkdir -p $OBJ
NVCC=${NVCC:-nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -O2 $ZKB_EXTRA_NVCC_FLAGS"
pids=()
for src in *.cu; do
  o=$OBJ/${src%.cu}.o
  newest=$(ls -t $src *.cuh ../../include/zkb.h | head -1)
  if [ ! -f $o ] || [ $newest -nt $o ]; then
    echo "nvcc $src"
    ( $NVCC $FLAGS -c $src -o $o ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
$NVCC -shared -o $OUT $OBJ/*.o -gencode arch=compute_100a,code=sm_100a
echo "built $(realpath $OUT)"
