#!/bin/bash
# Second multi-GPU pass: torchrun correctness with full log, push algorithm vs the others, NVLS / push launch-shape tuning.
N=${1:-8}
OUT=gpurun_out/t$N
mkdir -p $OUT
(timeout 300 python -m pytest tests/test_multigpu.py -m gpu -x -q -k torchrun 2>&1 | tail -5) > $OUT/pytest.log
tail -3 $OUT/pytest.log
grep -E "MISMATCH|Error|rror:|PASSED|FAILED|world=|Traceback|B2RError" gpurun_out/dist_worker_$N.log | sort | uniq -c | sort -rn | head -20
timeout 200 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=reduce,allreduce --algos=push,twoshot,nvls --min-bytes=1048576 --max-bytes=1073741824 --factor=4 --check > $OUT/push_sweep.txt 2>&1
grep -E "push|PASSED|FAILED|VERIFY" $OUT/push_sweep.txt
echo "== NVLS tuning"
for sz in 268435456 1073741824; do for u in 4 8; do for c in 8 16 32 64; do for t in 128 256 512; do
  echo -n "nvls bytes $sz unroll $u ctas $c threads $t: "
  B2R_COLL_UNROLL=$u B2R_COLL_CTAS=$c B2R_COLL_THREADS=$t timeout 60 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=allreduce --algos=nvls --min-bytes=$sz --max-bytes=$sz --iters=12 2>&1 | grep -E "allreduce" | awk '{print $7, $9}'
done; done; done; done > $OUT/tune_nvls.txt 2>&1
sort -k12 -n -r $OUT/tune_nvls.txt | head -8
echo "== push tuning"
for c in 16 32 64 128; do for t in 256 512; do
  echo -n "push ctas $c threads $t: "
  B2R_PUSH_CTAS=$c B2R_COLL_THREADS=$t timeout 60 bin/bench_collective --gpus=$N --sweep --types=FLOAT --colls=reduce --algos=push --min-bytes=268435456 --max-bytes=268435456 --iters=12 2>&1 | grep -E "reduce " | awk '{print $7, $8}'
done; done > $OUT/tune_push.txt 2>&1
cat $OUT/tune_push.txt
echo done
