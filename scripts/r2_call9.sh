#!/bin/bash
O=gpurun_out/r2c9; mkdir -p $O
timeout 200 bin/bench_collective --gpus=2 --sweep --types=FLOAT --ops=SUM,MIN --colls=reduce,allreduce --algos=auto,push,twoshot --min-bytes=1048576 --max-bytes=1073741824 --factor=4 --check 2>&1 | tee $O/push2.txt | grep -v "^#" 
timeout 100 bin/b2r_mpirun -np 2 bin/reduce_gpu_mpi 2048 2 2>&1 | tail -8
