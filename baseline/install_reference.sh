#!/bin/bash
# "Installs" the reference into baseline/_ref (git-ignored). The reference is not a Python package, so the pip route
# of the task statement fails (recorded in DESIGN.md); what can be installed is its source, copied verbatim:
#   - cuda/C/src/reduction/{reduction.cpp,reduction_kernel.cu,reduction.h}   -> baseline/_ref/src/
#   - mpi/{reduce.c,externalfunctions.h,constants.h}                          -> baseline/_ref/mpi/
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF=${REFERENCE_ROOT:-/root/reference}
[ -d "$REF" ] || { echo "reference tree $REF not found" >&2; exit 1; }
mkdir -p "$HERE/_ref/src" "$HERE/_ref/mpi"
cp "$REF"/cuda/C/src/reduction/reduction.cpp "$REF"/cuda/C/src/reduction/reduction_kernel.cu "$REF"/cuda/C/src/reduction/reduction.h "$HERE/_ref/src/"
cp "$REF"/mpi/reduce.c "$REF"/mpi/externalfunctions.h "$REF"/mpi/constants.h "$HERE/_ref/mpi/"
( cd "$HERE/_ref" && sha256sum src/* mpi/* > MANIFEST.sha256 )
echo "installed reference sources into $HERE/_ref"
