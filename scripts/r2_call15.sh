#!/bin/bash
N=${1:-2}; O=gpurun_out/r2c15_$N; mkdir -p $O
timeout 300 bin/bench_collective --gpus=$N --max-bytes=1073741824 --check --check-max=67108864 --timeout-ms=4000 --iters=10 --plan=scripts/plans/plan_hybrid.txt --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err; tail -2 $O/plan.txt; grep -c "VERIFY FAILED\|skipped or failed" $O/plan.txt; tail -3 $O/plan.err
timeout 100 bin/nccl_bench --gpus=$N --sweep --types=FLOAT --ops=SUM --colls=reduce --min-bytes=16777216 --max-bytes=1073741824 --factor=4 --csv=$O/nccl.csv 2>/dev/null | grep reduce
for s in 16777216 67108864 268435456 1073741824; do echo "== $s"; grep " FLOAT " $O/plan.txt | grep " $s " | sort -k7 -n | head -4 | cut -c1-170; done
grep "INT\|DOUBLE\|BF16" $O/plan.txt | cut -c1-140
[ "$2" = "tests" ] && timeout 400 python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -4
