#!/bin/bash
# round 2, record call (N GPUs): the reference's loop (ours / our tree / NCCL), BASELINE config 5 sweep (AUTO vs NCCL), D3, bench.py
N=${1:-8}
O=gpurun_out/r2c13_$N; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
timeout 200 bin/bench_collective --gpus=$N --check --jsonl=$O/refloop.jsonl --raw=$O/refloop_b2r.txt > $O/refloop_b2r.log 2>&1; tail -3 $O/refloop_b2r.log
timeout 300 bin/bench_collective --gpus=$N --algo=tree --retries=2 --check --raw=$O/refloop_tree.txt > $O/refloop_tree.log 2>&1; tail -2 $O/refloop_tree.log
timeout 200 bin/nccl_bench --gpus=$N --jsonl=$O/refloop.jsonl > $O/refloop_nccl.txt 2>/dev/null; tail -2 $O/refloop_nccl.txt
timeout 400 bin/bench_collective --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=reduce,allreduce --algos=auto --min-bytes=1024 --max-bytes=4294967296 --factor=4 --check --csv=$O/sweep.csv --jsonl=$O/sweep.jsonl > $O/sweep_b2r.txt 2>&1; tail -3 $O/sweep_b2r.txt
timeout 400 bin/nccl_bench --gpus=$N --sweep --types=INT,DOUBLE,BF16 --colls=reduce,allreduce --min-bytes=1024 --max-bytes=4294967296 --factor=4 --csv=$O/sweep.csv --jsonl=$O/sweep.jsonl > $O/sweep_nccl.txt 2>/dev/null; tail -2 $O/sweep_nccl.txt
timeout 200 bin/bench_collective --gpus=$N --sweep --types=FLOAT --ops=SUM,MIN --colls=reduce,allreduce --algos=auto --min-bytes=16384 --max-bytes=1073741824 --factor=16 --check --csv=$O/sweep_f32.csv > $O/sweep_f32_b2r.txt 2>&1
timeout 200 bin/nccl_bench --gpus=$N --sweep --types=FLOAT --ops=SUM,MIN --colls=reduce,allreduce --min-bytes=16384 --max-bytes=1073741824 --factor=16 --csv=$O/sweep_f32.csv > $O/sweep_f32_nccl.txt 2>/dev/null
timeout 200 bin/b2r_mpirun -np $N bin/reduce_gpu_mpi > $O/reduce_gpu_mpi.txt 2>&1; tail -2 $O/reduce_gpu_mpi.txt
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
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
python scripts/sweep_table.py $O/sweep.csv | awk -F'|' 'NR>2{printf "%-10s %-7s %-9s %-8s %10s %10s %s\n",$2,$3,$6,$7,$8,$10,$12}'
