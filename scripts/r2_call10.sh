#!/bin/bash
# round 2, call 10 (N GPUs, default 8): push vs AUTO, crossovers, then the record sweeps with those crossovers exported
N=${1:-8}
O=gpurun_out/r2c10_$N; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
C="--gpus=$N --max-bytes=1073741824 --check --check-max=16777216 --timeout-ms=4000 --iters=20"
timeout 300 bin/bench_collective $C --plan=scripts/plans/plan8_push.txt --csv=$O/plan.csv > $O/plan_push.txt 2> $O/plan_push.err; tail -4 $O/plan_push.txt; grep -c "VERIFY FAILED\|skipped or failed" $O/plan_push.txt
eval "$(python scripts/pick_push_crossovers.py $O/plan.csv)"; env | grep B2R_AUTO_PUSH | tee $O/crossovers.txt
# ---- the reference's loop (mpi/reduce.c: 2 GiB per datatype sharded over the ranks, 5 x {MAX, MIN, SUM} x {INT, DOUBLE}): ours, our tree, NCCL
timeout 200 bin/bench_collective --gpus=$N --check --jsonl=$O/refloop.jsonl --raw=$O/refloop_b2r.txt > $O/refloop_b2r.log 2>&1; tail -4 $O/refloop_b2r.log
timeout 300 bin/bench_collective --gpus=$N --algo=tree --retries=2 --check --raw=$O/refloop_tree.txt > $O/refloop_tree.log 2>&1; tail -3 $O/refloop_tree.log
timeout 200 bin/nccl_bench --gpus=$N --jsonl=$O/refloop.jsonl > $O/refloop_nccl.txt 2>/dev/null; tail -3 $O/refloop_nccl.txt
# ---- BASELINE config 5: message-size sweep 1 KB - 4 GB, int32 / fp64 / bf16, reduce + allreduce, AUTO vs NCCL, NVLink bytes measured
timeout 400 bin/bench_collective --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=reduce,allreduce --algos=auto --min-bytes=1024 --max-bytes=4294967296 --factor=4 --nvlink --check --csv=$O/sweep.csv --jsonl=$O/sweep.jsonl > $O/sweep_b2r.txt 2>&1; tail -3 $O/sweep_b2r.txt
timeout 400 bin/nccl_bench --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=reduce,allreduce --min-bytes=1024 --max-bytes=4294967296 --factor=4 --csv=$O/sweep.csv --jsonl=$O/sweep.jsonl > $O/sweep_nccl.txt 2>/dev/null; tail -2 $O/sweep_nccl.txt
# ---- D3: the reduce.c loop launched by (mini-)MPI, one rank per GPU, data never leaves the GPUs
timeout 200 bin/b2r_mpirun -np $N bin/reduce_gpu_mpi > $O/reduce_gpu_mpi.txt 2>&1; tail -4 $O/reduce_gpu_mpi.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 300 $T --master-port 29601 scripts/staged_vs_symm.py > $O/staged_vs_symm.txt 2>$O/staged.err; grep -v "^\*\|OMP\|^$" $O/staged_vs_symm.txt | tail -10
timeout 500 $T --master-port 29602 bench.py --gpus $N --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; tail -c 300 $O/bench.err
python - <<PY
import json
try:
    d=json.loads([l for l in open('$O/bench.json') if l.startswith('{')][-1])
    print('value',round(d['value']),'ms/step',d['ms_per_step'],'e2e',round(d['e2e']['value'],1))
    for r in d.get('collectives',{}).get('rows',[]): print(r['coll'],r['dtype'],r['bytes'],r['algo'],round(r['b2r_busbw_gbs'],1),round(r['nccl_busbw_gbs'],1),round(r['ratio'],3),r['b2r_check'],r['nccl_check'])
    print(d.get('collectives',{}).get('error'), {k:v for k,v in d.get('config4',{}).items() if k in ('b2r_gbs','library_gbs','ratio','self_check','error')})
except Exception as e: print('parse failed',e)
PY

[ "$2" = "tests" ] && timeout 400 python -m pytest tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -4
