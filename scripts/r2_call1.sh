#!/bin/bash
# round 2, call 1 (1 GPU): experimental kernels in loopback, sanitizers
mkdir -p gpurun_out/r2c1
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,clocks_throttle_reasons.active --format=csv > gpurun_out/r2c1/smi.txt 2>&1
B2R_TEST_EXPERIMENTAL=1 timeout 900 python -m pytest tests/test_experimental_gpu.py -m gpu -q 2>&1 | tail -40 > gpurun_out/r2c1/experimental.txt
for t in memcheck racecheck synccheck; do timeout 600 bash scripts/sanitize.sh $t > gpurun_out/r2c1/sanitize_$t.tail 2>&1; done
tail -5 gpurun_out/r2c1/experimental.txt; tail -3 gpurun_out/r2c1/sanitize_*.tail
