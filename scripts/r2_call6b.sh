#!/bin/bash
# 2-GPU dry run of the 8-GPU plan + re-check after the relaxed-ack change
O=gpurun_out/r2c6b; mkdir -p $O
C="--gpus=2 --max-bytes=1073741824 --check --check-max=16777216 --timeout-ms=4000 --iters=10"
timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_a.txt > $O/plan_a.txt 2> $O/plan_a.err; tail -3 $O/plan_a.txt; tail -3 $O/plan_a.err
timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_b.txt > $O/plan_b.txt 2> $O/plan_b.err; tail -3 $O/plan_b.txt; tail -3 $O/plan_b.err
B2R_LL_BIG_MB=16 timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_c.txt > $O/plan_c.txt 2> $O/plan_c.err; tail -3 $O/plan_c.txt; tail -3 $O/plan_c.err
grep -c "VERIFY FAILED" $O/plan_*.txt; grep -h "FAILED\|skipped or failed" $O/plan_*.txt | head
echo "== oneshot unroll A/B"; for u in 4 8; do B2R_COLL_UNROLL=$u bin/bench_collective --gpus=2 --sweep --types=FLOAT,INT --colls=reduce --algos=oneshot --min-bytes=67108864 --max-bytes=67108864 2>/dev/null | grep reduce; done
echo "== trace"; python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/xg_trace.py --steps 200 2>&1 | grep -v "^\*\|OMP\|^$"
for i in 1 2 3; do python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2956$i scripts/xg_trace.py --steps 20 2>&1 | grep "us per step"; done
timeout 600 python -m pytest tests/test_zz_no_alloc.py tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -6
