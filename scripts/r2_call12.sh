#!/bin/bash
N=${1:-4}; O=gpurun_out/r2c12_$N; mkdir -p $O
timeout 300 bin/bench_collective --gpus=$N --max-bytes=1073741824 --timeout-ms=4000 --iters=10 --plan=scripts/plans/plan_int_nvls.txt --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err; tail -3 $O/plan.txt
for c in reduce allreduce; do for s in 16777216 67108864 1073741824; do echo "== $c $s"; grep " $c " $O/plan.txt | grep " $s " | sort -k7 -n | head -4 | cut -c1-170; done; done
