#!/bin/bash
# compute-sanitizer targets (the reference has none — its emulation-mode bank checker, cutil bank_checker.h,
# died with CUDA 3.1). ONE tool per gpurun call, on the smallest case that shows what is needed (B200_PROFILING.md):
#   gpurun -- bash scripts/sanitize.sh memcheck     # out-of-bounds / misaligned accesses in the single-GPU family
#   gpurun -- bash scripts/sanitize.sh racecheck    # shared-memory hazards (block reduce, TMA ring, pointer tables)
#   gpurun -- bash scripts/sanitize.sh synccheck    # barrier misuse (mbarrier, bar.sync, cluster barrier)
# The cross-rank protocol is exercised through the LOOPBACK transport (virtual ranks in one cooperative launch), which is
# the only multi-rank configuration a sanitizer can follow on one GPU.
set -e
TOOL=${1:-memcheck}
OUT=gpurun_out/sanitize_$TOOL.log
mkdir -p gpurun_out
CMD="python -m pytest tests/test_reduce_local_gpu.py -m gpu -x -q -k 'misaligned or tma_single or int_extremes'"
[ "$TOOL" = "racecheck" ] && CMD="python -m pytest tests/test_collectives_loopback_gpu.py -m gpu -x -q -k 'barrier or symmetric'"
eval "$CMD" > gpurun_out/sanitize_plain.log 2>&1   # must pass without the tool first
eval "timeout 900 compute-sanitizer --tool $TOOL --error-exitcode 7 $CMD" > $OUT 2>&1
tail -15 $OUT
