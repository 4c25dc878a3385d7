#!/bin/bash
# fp32 / fp64 min and max through the switch as integer keys (keyed NVLS) against two-shot by peer loads, N GPUs, checked
N=${1:-4}
O=gpurun_out/r2c16_$N; mkdir -p $O
timeout 80 bin/bench_collective --gpus=$N --plan=scripts/plans/plan_keyed.txt --max-bytes=536870912 --check --check-max=268435456 --iters=6 --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err
grep -h "K\.\|FAILED\|clocks gpu0" $O/plan.txt $O/plan.err | cut -c1-120
