#!/bin/bash
# Record run on N GPUs (one invocation per tool: communicator set-up dominates short runs).
N=${1:-8}
OUT=gpurun_out/rec$N
mkdir -p $OUT
rm -f $OUT/*.csv $OUT/*.jsonl
(timeout 200 python -m pytest tests/test_multigpu.py -m gpu -x -q -k torchrun 2>&1 | tail -3) > $OUT/pytest.log; tail -2 $OUT/pytest.log
grep -c MISMATCH gpurun_out/dist_worker_$N.log
timeout 150 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=reduce,allreduce --algos=auto --min-bytes=1024 --max-bytes=4294967296 --factor=4 --check --csv=$OUT/sweep.csv --jsonl=$OUT/sweep.jsonl > $OUT/b2r_float.txt 2>&1; tail -2 $OUT/b2r_float.txt
timeout 100 bin/bench_collective --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=reduce,allreduce --algos=auto --min-bytes=4096 --max-bytes=1073741824 --factor=16 --check --csv=$OUT/sweep.csv --jsonl=$OUT/sweep.jsonl > $OUT/b2r_types.txt 2>&1; tail -1 $OUT/b2r_types.txt
timeout 100 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=reduce,allreduce --algos=oneshot,twoshot,tree,push,nvls --min-bytes=65536 --max-bytes=1073741824 --factor=128 --csv=$OUT/sweep_algos.csv > $OUT/b2r_algos.txt 2>&1
timeout 100 bin/bench_collective --gpus=$N --sweep --types=FLOAT,INT --ops=MIN --colls=reduce,allreduce --algos=auto --min-bytes=1048576 --max-bytes=268435456 --factor=16 --check --csv=$OUT/sweep.csv > $OUT/b2r_min.txt 2>&1
timeout 120 bin/nccl_bench --gpus=$N --sweep --types=FLOAT --min-bytes=1024 --max-bytes=4294967296 --factor=4 --csv=$OUT/sweep.csv --jsonl=$OUT/sweep.jsonl > $OUT/nccl_float.txt 2>/dev/null
timeout 100 bin/nccl_bench --gpus=$N --sweep --types=INT,DOUBLE,BF16 --min-bytes=4096 --max-bytes=1073741824 --factor=16 --csv=$OUT/sweep.csv --jsonl=$OUT/sweep.jsonl > $OUT/nccl_types.txt 2>/dev/null
timeout 60 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT --ops=MIN --min-bytes=1048576 --max-bytes=268435456 --factor=16 --csv=$OUT/sweep.csv > $OUT/nccl_min.txt 2>/dev/null
timeout 100 bin/bench_collective --gpus=$N --check --jsonl=$OUT/refloop.jsonl --raw=$OUT/refloop_b2r.txt > $OUT/refloop_b2r.log 2>&1; tail -1 $OUT/refloop_b2r.log
timeout 100 bin/nccl_bench --gpus=$N --jsonl=$OUT/refloop.jsonl > $OUT/refloop_nccl.txt 2>/dev/null
timeout 150 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 2000 --warmup 50 2>&1 | grep metric > $OUT/bench.json
cut -c1-220 $OUT/bench.json
echo done
