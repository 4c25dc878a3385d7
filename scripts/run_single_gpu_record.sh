#!/bin/bash
# Single-GPU record: the reference sample's configurations (2^24 elements, {int,float,double,bf16} x {SUM,MIN,MAX}) in plain
# stream order and with PDL, the size sweep 2^10..2^30 per dtype (BASELINE.json config 3), and cub::DeviceReduce.
OUT=$PWD/gpurun_out/single
mkdir -p $OUT; rm -f $OUT/*.jsonl $OUT/*.csv
cd bin
for t in int float double bf16; do for m in SUM MIN MAX; do
  ./bench_reduce --method=$m --type=$t --nolog --iters=200 --jsonl=$OUT/plain.jsonl | grep -E "Device-event|FAILED" | sed "s/^/$t $m plain: /"
  ./bench_reduce --method=$m --type=$t --nolog --iters=200 --pdl --jsonl=$OUT/pdl.jsonl | grep -E "Device-event|FAILED" | sed "s/^/$t $m pdl:   /"
done; done > $OUT/sample_configs.txt 2>&1
for t in int float double bf16; do ./bench_reduce --method=SUM --type=$t --shmoo --n=1073741824 --iters=50 --nolog --csv=$OUT/shmoo.csv; done > $OUT/shmoo.txt 2>&1
./bench_reduce --method=MAX --type=double --shmoo --n=268435456 --iters=50 --nolog --csv=$OUT/shmoo.csv >> $OUT/shmoo.txt 2>&1
for t in int float double bf16; do for m in SUM MAX; do ./cub_bench --type=$t --method=$m --max=1073741824 --csv=$OUT/cub.csv; done; done > $OUT/cub.txt 2>&1
for t in int double; do ../baseline/_ref/bin/reduction --method=SUM --type=$t --noprompt | grep Throughput | sed "s/^/reference $t SUM: /"; ../baseline/_ref/bin/reduction --method=MAX --type=$t --noprompt | grep -E "Throughput|PASSED|FAILED" | tr '\n' ' ' | sed "s/^/reference $t MAX: /"; echo; done > $OUT/reference_sample.txt 2>&1
rm -f reduction.txt SdkMasterLog.csv
tail -3 $OUT/sample_configs.txt; tail -4 $OUT/reference_sample.txt
