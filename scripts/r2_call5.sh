#!/bin/bash
# round 2, call 5 (2 GPUs): multi-GPU tests, bench.py --gpus 2 (collectives + config4 sections), reference arm, single-GPU A/B
O=gpurun_out/r2c5; mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
echo "== A/B packed sum, early PDL (1 GPU)"
for p in 1 0; do B2R_PACKED_SUM=$p timeout 120 python bench.py --steps 20 --warmup 5 2>$O/ab_$p.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('packed=$p', round(d['value']), round(d['value_plain_stream_order']), round(d['value_no_programmatic_launch']), round(d['value_host_synced']), d['clocks']['samples'], d['clocks']['sm_mhz'])"; done
B2R_PACKED_SUM=1 timeout 100 bin/local_sweep --pdl=2 --iters=300 --quick 2>&1 | head -6
B2R_PACKED_SUM=0 timeout 100 bin/local_sweep --pdl=2 --iters=300 --quick 2>&1 | head -6
echo "== other dtypes, default plain geometry candidates"
for t in double float bf16; do n=16777216; [ $t = double ] && n=8388608; [ $t = bf16 ] && n=33554432; timeout 100 bin/local_sweep --pdl=1 --dyn --type=$t --n=$n --iters=200 2>&1 | sed -n 3,5p; done
echo "== multi-GPU tests"
timeout 900 python -m pytest tests/test_multigpu.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -15 > $O/pytest.txt; tail -4 $O/pytest.txt
grep -i "mismatch\|error" gpurun_out/dist_worker_2.log | head
echo "== bench 2 GPUs"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 20 --warmup 5 > $O/bench2.json 2> $O/bench2.err
tail -c 400 $O/bench2.err
python - <<'PY'
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2c5/bench2.json') if l.startswith('{')][-1])
    print('value',round(d['value']),'ms/step',d['ms_per_step'],'e2e',round(d['e2e']['value'],1),'clocks',d['clocks'])
    for r in d.get('collectives',{}).get('rows',[]): print(r['coll'],r['dtype'],r['bytes'],r['algo'],round(r['b2r_busbw_gbs'],1),round(r['nccl_busbw_gbs'],1),round(r['ratio'],3),r['b2r_check'],r['nccl_check'])
    print(d.get('collectives',{}).get('error'), d.get('config4'))
except Exception as e: print('parse failed',e)
PY
echo "== reference arm"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --impl reference --gpus 2 --steps 20 --warmup 5 > $O/ref2.json 2> $O/ref2.err; cut -c1-900 $O/ref2.json
timeout 200 python bench.py --impl reference --steps 20 --warmup 5 > $O/ref1.json 2>$O/ref1.err; cut -c1-300 $O/ref1.json
