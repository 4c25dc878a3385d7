#!/bin/bash
# posted reduce on 8 GPUs: posted vs the previous choice (plan, checked), NCCL on the same sizes, and the AUTO sweep as shipped
N=${1:-8}
O=gpurun_out/r2c12_$N; mkdir -p $O
timeout 150 bin/bench_collective --gpus=$N --plan=scripts/plans/plan_posted.txt --check --iters=10 --csv=$O/plan.csv > $O/plan.txt 2> $O/plan.err; grep -c "Q\." $O/plan.txt; grep -h "FAILED\|clocks gpu0" $O/plan.txt $O/plan.err | head -n 6
timeout 120 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT,DOUBLE,BF16 --ops=SUM,MIN,MAX --colls=reduce --min-bytes=524288 --max-bytes=67108864 --factor=2 --csv=$O/nccl.csv > $O/nccl.txt 2>/dev/null
timeout 120 bin/bench_collective --gpus=$N --sweep --types=INT,DOUBLE,BF16,FLOAT --ops=SUM --colls=reduce --algos=auto --min-bytes=1024 --max-bytes=67108864 --factor=4 --check --csv=$O/sweep.csv > $O/sweep_b2r.txt 2>&1; tail -n 2 $O/sweep_b2r.txt
timeout 120 bin/nccl_bench --gpus=$N --sweep --types=INT,DOUBLE,BF16,FLOAT --ops=SUM --colls=reduce --min-bytes=1024 --max-bytes=67108864 --factor=4 --csv=$O/sweep.csv > $O/sweep_nccl.txt 2>/dev/null
python scripts/posted_table.py $O/plan.csv $O/nccl.csv > $O/table.txt; grep -v "MIN   \|FLOAT  MAX\|INT    MIN" $O/table.txt
python scripts/sweep_table.py $O/sweep.csv | awk -F'|' 'NR>2{printf "%-10s %-7s %-9s %-8s %10s %10s %s\n",$2,$3,$6,$7,$8,$10,$12}'
