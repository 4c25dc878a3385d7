#!/bin/bash
# posted reduce on N real GPUs: the multimem path checked (torchrun worker, short form), then posted vs the previous choice vs NCCL
N=${1:-4}
O=gpurun_out/r2c11_$N; mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
B2R_DIST_ONLY=posted timeout 200 $T --master-port 29611 tests/dist_worker.py > $O/dist_posted.txt 2>&1; grep -h "DIST_WORKER\|MISMATCH\|multicast" $O/dist_posted.txt | head -n 12
timeout 300 bin/bench_collective --gpus=$N --plan=scripts/plans/plan_posted.txt --check --iters=10 --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err; grep -c "Q\." $O/plan.txt; grep -h "FAILED\|clocks" $O/plan.txt | head -n 6
timeout 200 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT,DOUBLE,BF16 --ops=SUM --colls=reduce --min-bytes=524288 --max-bytes=67108864 --factor=2 --csv=$O/nccl.csv > $O/nccl_sum.txt 2>/dev/null
timeout 200 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,DOUBLE,INT --ops=MIN,MAX --colls=reduce --min-bytes=524288 --max-bytes=67108864 --factor=2 --csv=$O/nccl.csv > $O/nccl_minmax.txt 2>/dev/null
python scripts/posted_table.py $O/plan.csv $O/nccl.csv
