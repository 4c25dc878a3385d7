#!/bin/bash
# round 2, call 6 (2 GPUs): per-phase trace N=1 vs N=2, native harness sanity (clocks line, NVLink counters), 2-GPU oddities
O=gpurun_out/r2c6; mkdir -p $O
python scripts/xg_trace.py --steps 200 > $O/trace_n1.txt 2>&1; cat $O/trace_n1.txt
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 scripts/xg_trace.py --steps 200 > $O/trace_n2.txt 2>&1; grep -v "^\*\|OMP\|^$" $O/trace_n2.txt
echo "== 20-step runs"; for i in 1 2 3; do python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2956$i scripts/xg_trace.py --steps 20 2>&1 | grep "us per step"; done
for i in 1 2 3; do python scripts/xg_trace.py --steps 20 | grep "us per step"; done
echo "== harness"; timeout 200 bin/bench_collective --gpus=2 --sweep --types=FLOAT,INT --colls=reduce,allreduce --algos=auto,oneshot,twoshot,ring --min-bytes=67108864 --max-bytes=67108864 --nvlink --check 2>&1 | tee $O/sweep2.txt
B2R_RING_V2=1 timeout 100 bin/bench_collective --gpus=2 --sweep --types=FLOAT --colls=reduce --algos=ring --min-bytes=1048576 --max-bytes=268435456 --factor=16 --check 2>&1 | tee $O/ring2.txt
timeout 100 bin/bench_collective --gpus=2 --sweep --types=FLOAT,INT --colls=reduce,allreduce --algos=auto,ll --min-bytes=1024 --max-bytes=65536 --factor=8 --check 2>&1 | tee $O/small.txt
timeout 100 bin/nccl_bench --gpus=2 --sweep --types=FLOAT --colls=reduce,allreduce --min-bytes=1024 --max-bytes=67108864 --factor=16 2>/dev/null | tee $O/nccl.txt
timeout 600 python -m pytest tests/test_zz_no_alloc.py tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -4
timeout 60 bin/bench_reduce --method=SUM --type=int --noprompt | grep -i "clocks\|Throughput\|PASS"
