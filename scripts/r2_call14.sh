#!/bin/bash
# posted reduce with coalesced 4- / 8-byte multimem requests, N GPUs: checked rows only (the other arms were measured in r2_call12)
N=${1:-8}
O=gpurun_out/r2c14_$N; mkdir -p $O
timeout 100 bin/bench_collective --gpus=$N --plan=scripts/plans/plan_posted_narrow.txt --check --iters=10 --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err
grep -h "R\.\|FAILED\|clocks gpu0" $O/plan.txt $O/plan.err | cut -c1-110
