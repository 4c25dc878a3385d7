#!/bin/bash
# Builds every baseline binary: the unmodified reference sample (sm_100), the cub::DeviceReduce bench, the NCCL bench.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
mkdir -p "$ROOT/bin"
[ -f "$HERE/_ref/src/reduction.cpp" ] || bash "$HERE/install_reference.sh" || true
[ -f "$HERE/_ref/src/reduction.cpp" ] && bash "$HERE/build_reference.sh"
$NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -w -I"$ROOT/include" "$HERE/cub_bench.cu" -o "$ROOT/bin/cub_bench" \
  -L"$ROOT/cuda_mpi_reductions_b200/lib" -lb2r -ldl -Xlinker -rpath,'$ORIGIN/../cuda_mpi_reductions_b200/lib'
if [ -f "$HERE/nccl_bench.cu" ]; then
  $NVCC -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -w -I"$ROOT/include" "$HERE/nccl_bench.cu" -o "$ROOT/bin/nccl_bench" \
    -L"$ROOT/cuda_mpi_reductions_b200/lib" -lb2r -lnccl -ldl -Xlinker -rpath,'$ORIGIN/../cuda_mpi_reductions_b200/lib'
fi
echo "baselines built"
