#!/bin/bash
# Multi-GPU measurement pass (run under `gpurun --gpus N`): correctness at N ranks, collective sweeps (ours vs NCCL),
# launch-shape tuning at 256 MiB, bench.py under torchrun, and the reference-format reduce.c loop.
N=${1:-8}
OUT=gpurun_out/r$N
mkdir -p $OUT
nvidia-smi topo -m > $OUT/topo.txt 2>&1
(timeout 420 python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -30) > $OUT/pytest.log
tail -4 $OUT/pytest.log
timeout 300 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=allreduce,reduce --algos=auto,oneshot,twoshot,nvls --min-bytes=1024 --max-bytes=1073741824 --factor=4 --check --csv=$OUT/sweep.csv > $OUT/sweep_float.txt 2>&1
tail -3 $OUT/sweep_float.txt
NCCL_DEBUG=INFO timeout 200 bin/nccl_bench --gpus=$N --sweep --types=FLOAT --min-bytes=1024 --max-bytes=1073741824 --factor=4 --csv=$OUT/sweep.csv > $OUT/nccl_float.txt 2> $OUT/nccl_debug.txt
grep -i -m3 "nvls" $OUT/nccl_debug.txt
for u in 4 8; do for c in 32 64 128 148; do for t in 256 512; do
  echo "== unroll $u ctas $c threads $t"
  B2R_COLL_UNROLL=$u B2R_COLL_CTAS=$c B2R_COLL_THREADS=$t timeout 60 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=allreduce,reduce --algos=twoshot,nvls --min-bytes=268435456 --max-bytes=268435456 2>&1 | grep -E "allreduce|reduce "
done; done; done > $OUT/tune_256m.txt 2>&1
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 1000 --warmup 20 2>&1 | grep metric > $OUT/bench.json
cut -c1-200 $OUT/bench.json
timeout 120 bin/bench_collective --gpus=$N --jsonl=$OUT/refloop.jsonl --raw=$OUT/refloop_b2r.txt > /dev/null 2>&1
timeout 120 bin/nccl_bench --gpus=$N --jsonl=$OUT/refloop.jsonl > $OUT/refloop_nccl.txt 2>/dev/null
timeout 200 bin/bench_collective --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=allreduce,reduce --algos=auto --min-bytes=1048576 --max-bytes=1073741824 --factor=32 --csv=$OUT/sweep.csv > $OUT/sweep_types.txt 2>&1
timeout 200 bin/nccl_bench --gpus=$N --sweep --types=INT,DOUBLE,BF16 --min-bytes=1048576 --max-bytes=1073741824 --factor=32 --csv=$OUT/sweep.csv > $OUT/nccl_types.txt 2>/dev/null
echo done
