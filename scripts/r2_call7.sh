#!/bin/bash
# round 2, call 7 (8 GPUs): every candidate algorithm / launch shape for reduce and allreduce in three processes, NCCL on the
# same sizes, then bench.py (headline + collectives + config 4) and the per-phase trace at 8 GPUs.
O=gpurun_out/r2c7; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
C="--gpus=8 --max-bytes=1073741824 --check --check-max=16777216 --timeout-ms=4000 --iters=20"
timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_a.txt --csv=$O/plan.csv > $O/plan_a.txt 2> $O/plan_a.err; tail -3 $O/plan_a.txt
timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_b.txt --csv=$O/plan.csv > $O/plan_b.txt 2> $O/plan_b.err; tail -3 $O/plan_b.txt
B2R_LL_BIG_MB=16 timeout 400 bin/bench_collective $C --plan=scripts/plans/plan8_c.txt --csv=$O/plan.csv > $O/plan_c.txt 2> $O/plan_c.err; tail -3 $O/plan_c.txt
timeout 300 bin/nccl_bench --gpus=8 --sweep --types=FLOAT,INT,DOUBLE,BF16 --ops=SUM --colls=reduce,allreduce --min-bytes=1024 --max-bytes=1073741824 --factor=4 --csv=$O/nccl.csv > $O/nccl_sum.txt 2>/dev/null; tail -2 $O/nccl_sum.txt
timeout 200 bin/nccl_bench --gpus=8 --sweep --types=FLOAT,INT --ops=MIN --colls=reduce,allreduce --min-bytes=1024 --max-bytes=1073741824 --factor=4 --csv=$O/nccl.csv > $O/nccl_min.txt 2>/dev/null; tail -2 $O/nccl_min.txt
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 bench.py --gpus 8 --steps 20 --warmup 5 > $O/bench8.json 2> $O/bench8.err; tail -c 300 $O/bench8.err; cut -c1-400 $O/bench8.json
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29572 scripts/xg_trace.py --steps 200 > $O/trace_n8.txt 2>&1; grep "us per step" $O/trace_n8.txt
for i in 1 2; do timeout 100 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2958$i scripts/xg_trace.py --steps 20 2>&1 | grep "us per step" | head -3; done
grep -c FAILED $O/plan_*.txt; grep -h "FAILED\|skipped or failed" $O/plan_*.txt | head -20
B2R_BENCH_GRAPH=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29573 bench.py --gpus 8 --steps 20 --warmup 5 --no-collectives 2>/dev/null | grep '^{' | cut -c1-260
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29574 bench.py --gpus 8 --steps 20 --warmup 5 --no-collectives 2>/dev/null | grep '^{' | cut -c1-260
