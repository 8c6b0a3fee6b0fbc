#!/usr/bin/env bash
# Builds libgpk.so in-tree for B200 (sm_100a).  Usage: build.sh [extra nvcc flags]
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
OUT=../libgpk.so
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
mkdir -p build
objs=()
for f in gpk_api gpk_cov gpk_diag gpk_gemm gpk_kmeans gpk_loss gpk_oz gpk_pso gpk_surface; do
  if [ ! -f build/$f.o ] || [ $f.cu -nt build/$f.o ] || [ gpk_common.cuh -nt build/$f.o ] || [ gpk_cholinv.cuh -nt build/$f.o ] || [ gpk_ozdig.cuh -nt build/$f.o ] || [ ../../include/gpk.h -nt build/$f.o ]; then
    $NVCC $FLAGS "$@" -c $f.cu -o build/$f.o &
  fi
  objs+=(build/$f.o)
done
for f in gpk_csv gpk_finish; do
  if [ ! -f build/$f.o ] || [ $f.cpp -nt build/$f.o ] || [ ../../include/gpk.h -nt build/$f.o ]; then
    $NVCC $FLAGS "$@" -x cu -c $f.cpp -o build/$f.o &
  fi
  objs+=(build/$f.o)
done
wait
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o $OUT "${objs[@]}"
echo "built $(readlink -f $OUT)"
