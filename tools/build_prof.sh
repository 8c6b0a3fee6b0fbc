#!/usr/bin/env bash
# Instrumented libgpk (-DOZ_PROFILE: cycle counters in the MMA thread of the digit GEMM) -> tools/libgpk_prof.so
set -euo pipefail
cd "$(dirname "$0")/../gaussian-process_b200/csrc"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden -DOZ_PROFILE ${EXTRA:-}"
mkdir -p build/prof
for f in gpk_api gpk_cov gpk_diag gpk_gemm gpk_loss gpk_oz gpk_surface; do
  $NVCC $FLAGS -c $f.cu -o build/prof/$f.o &
done
$NVCC $FLAGS -x cu -c gpk_csv.cpp -o build/prof/gpk_csv.o &
$NVCC $FLAGS -x cu -c gpk_finish.cpp -o build/prof/gpk_finish.o &
wait
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o ../../tools/libgpk_prof.so build/prof/*.o
echo "built tools/libgpk_prof.so"
