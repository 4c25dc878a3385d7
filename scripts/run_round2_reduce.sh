#!/bin/bash
# First thing to run in the next round (one `gpurun --gpus N` call, N = 4 or 8): where does the store-forwarding ring
# stand against the distributed-NVLS reduce and NCCL, and which launch shape does it want? One process per tool
# (communicator set-up dominates short runs at 8 GPUs — the round-1 tuning loops burned 84 GPU-minutes on it).
N=${1:-8}
OUT=gpurun_out/round2_reduce_$N
mkdir -p $OUT
timeout 120 bin/bench_collective --gpus=$N --tune=ring --bytes=268435456 --iters=8 > $OUT/tune_ring_256m.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --tune=push --coll=reduce --bytes=268435456 --iters=8 > $OUT/tune_push_256m.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --sweep --types=FLOAT,INT --colls=reduce --algos=ring,nvls,twoshot --min-bytes=1048576 --max-bytes=1073741824 --factor=4 --check > $OUT/reduce_algos.txt 2>&1
timeout 100 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT --colls=reduce --min-bytes=1048576 --max-bytes=1073741824 --factor=4 > $OUT/nccl_reduce.txt 2>/dev/null
sort -k5 -n -r $OUT/tune_ring_256m.txt | head -5
grep -E "ring|FAIL|PASS" $OUT/reduce_algos.txt | tail -14
