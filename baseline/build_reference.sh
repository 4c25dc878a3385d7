#!/bin/bash
# Builds the UNMODIFIED reference reduction sample for sm_100 into baseline/_ref/bin/reduction.
#   sources : baseline/_ref/src/{reduction.cpp,reduction_kernel.cu,reduction.h}  (byte-for-byte copies of
#             /root/reference/cuda/C/src/reduction/*, fetched by baseline/install_reference.sh; baseline/_ref is
#             git-ignored but travels to the GPU box)
#   shim    : baseline/shim/*.h — our stand-in for cutil/shrUtils/shrQATest (the SDK's prebuilt .a files are 2012 objects)
# The kernels are built as-is, including their post-Volta-racy `volatile` warp-synchronous tail.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
SRC="$HERE/_ref/src"
OUT="$HERE/_ref/bin"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
[ -f "$SRC/reduction.cpp" ] || { echo "missing $SRC/reduction.cpp — run baseline/install_reference.sh first" >&2; exit 1; }
mkdir -p "$OUT"
# -arch=sm_100 (not sm_100a): nothing in the 2012 source needs arch-specific features. math_functions.h is found in the toolkit.
$NVCC -arch=sm_100 -O2 -Xcompiler -fno-strict-aliasing -w -I"$HERE/shim" -I"$SRC" -x cu "$SRC/reduction.cpp" "$SRC/reduction_kernel.cu" -o "$OUT/reduction"
echo "built $OUT/reduction"
# end-to-end driver around the SAME unmodified kernel file (see ref_e2e.cu): per-step H2D + two-pass reduce + D2H
$NVCC -arch=sm_100 -O2 -Xcompiler -fno-strict-aliasing -w -I"$SRC" "$HERE/ref_e2e.cu" "$SRC/reduction_kernel.cu" -o "$OUT/ref_e2e"
echo "built $OUT/ref_e2e"
