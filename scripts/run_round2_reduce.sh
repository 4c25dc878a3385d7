#!/bin/bash
# First things to run in the next round.
#   scripts/run_round2_reduce.sh validate      one GPU: the experimental kernels written blind at the end of round 1 (LL small-message path,
#                                              ring allreduce, pipelined ring) against their host restatements on the loopback transport
#   scripts/run_round2_reduce.sh N             one `gpurun --gpus N` call, N = 4 or 8: where do the store-forwarding ring, the LL path and the ring
#                                              allreduce stand against what AUTO picks today and against NCCL, and which launch shape does the ring want?
# One process per tool (communicator set-up dominates short runs at 8 GPUs — the round-1 tuning loops burned 84 GPU-minutes on it).
if [ "$1" = "validate" ]; then
  mkdir -p gpurun_out/round2_validate
  B2R_TEST_EXPERIMENTAL=1 timeout 1500 python -m pytest tests/test_experimental_gpu.py -m gpu -x -q 2>&1 | tee gpurun_out/round2_validate/pytest.txt | tail -15
  exit ${PIPESTATUS[0]}
fi
N=${1:-8}
OUT=gpurun_out/round2_reduce_$N
mkdir -p $OUT
export B2R_EXPERIMENTAL=1
timeout 120 bin/bench_collective --gpus=$N --tune=ring --bytes=268435456 --iters=8 > $OUT/tune_ring_256m.txt 2>&1
B2R_RING_V2=1 timeout 120 bin/bench_collective --gpus=$N --tune=ring --bytes=268435456 --iters=8 > $OUT/tune_ring2_256m.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --tune=ring --coll=allreduce --type=INT --bytes=268435456 --iters=8 > $OUT/tune_ringall_256m.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --tune=push --coll=reduce --bytes=268435456 --iters=8 > $OUT/tune_push_256m.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --sweep --types=FLOAT,INT --colls=reduce --algos=ring,nvls,twoshot --min-bytes=1048576 --max-bytes=1073741824 --factor=4 --check > $OUT/reduce_algos.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --sweep --types=INT --ops=SUM,MIN --colls=allreduce --algos=ring,twoshot --min-bytes=1048576 --max-bytes=1073741824 --factor=4 --check > $OUT/allreduce_ring.txt 2>&1
# 32-bit integer multimem ops move 4 bytes per instruction: never measured against the P2P two-shot at scale
timeout 120 bin/bench_collective --gpus=$N --sweep --types=INT --ops=SUM,MAX --colls=reduce,allreduce --algos=nvls,twoshot --min-bytes=65536 --max-bytes=268435456 --factor=16 --check > $OUT/int_nvls.txt 2>&1
timeout 120 bin/bench_collective --gpus=$N --sweep --types=FLOAT,INT --colls=reduce,allreduce --algos=auto,ll --min-bytes=1024 --max-bytes=16384 --factor=2 --check > $OUT/small_ll.txt 2>&1
B2R_LL_BIG_MB=16 timeout 120 bin/bench_collective --gpus=$N --sweep --types=FLOAT,INT --colls=reduce,allreduce --algos=auto,ll --min-bytes=32768 --max-bytes=262144 --factor=2 --check > $OUT/mid_ll_big.txt 2>&1
timeout 100 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT --colls=reduce,allreduce --min-bytes=1024 --max-bytes=262144 --factor=2 > $OUT/nccl_small.txt 2>/dev/null
timeout 100 bin/nccl_bench --gpus=$N --sweep --types=FLOAT,INT --colls=reduce,allreduce --min-bytes=1048576 --max-bytes=1073741824 --factor=4 > $OUT/nccl_large.txt 2>/dev/null
sort -k5 -n -r $OUT/tune_ring_256m.txt | head -5
sort -k5 -n -r $OUT/tune_ring2_256m.txt | head -5
sort -k5 -n -r $OUT/tune_ringall_256m.txt | head -3
grep -E "ring|FAIL|PASS" $OUT/reduce_algos.txt $OUT/allreduce_ring.txt | tail -24
grep -E " ll |FAIL|PASS" $OUT/small_ll.txt $OUT/mid_ll_big.txt | tail -40
grep -E "nvls|twoshot|FAIL|PASS" $OUT/int_nvls.txt | tail -16
# BASELINE.json config 4: 1 GiB fp64 sharded over the GPUs, fused local reduce + NVLink combine vs torch reduction + ncclReduce of the scalars
PER=$((134217728 / N))
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus $N --dtype float64 --n $PER --steps 300 --warmup 20 > $OUT/config4_b2r.json 2>$OUT/config4_b2r.err
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus $N --dtype float64 --n $PER --steps 300 --warmup 20 --impl torch_nccl > $OUT/config4_torch_nccl.json 2>$OUT/config4_torch_nccl.err
tail -c 600 $OUT/config4_b2r.json; tail -c 400 $OUT/config4_torch_nccl.json
# per-size table, best algorithm of ours vs NCCL, and the crossover points for b2r_select_algo
python scripts/compare_sweeps.py $OUT/*.txt --md $OUT/comparison.md | tail -40
