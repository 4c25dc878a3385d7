#!/bin/bash
O=gpurun_out/r2c6c; mkdir -p $O
timeout 600 python -m pytest tests/test_zz_no_alloc.py tests/test_multigpu.py -m gpu -x -q 2>&1 | tail -6
grep -i "mismatch\|error\|Traceback" gpurun_out/dist_worker_2.log | head
for i in 1 2; do timeout 200 python bench.py --steps 20 --warmup 5 2>$O/b1.err | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('N=1', round(d['value']), d['ms_per_step'], round(d['value_plain_stream_order']), round(d['value_no_programmatic_launch']), round(d['value_host_synced']), d['config']['launch_mode'][:30], d['clocks']['samples'])"; done; tail -2 $O/b1.err
for i in 1 2; do timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2954$i bench.py --gpus 2 --steps 20 --warmup 5 --no-collectives 2> $O/b2.err | python -c "
import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('N=2', round(d['value']), d['ms_per_step'], round(d['e2e']['value'],1), d['config']['launch_mode'][:30])"; done; grep -v "^\*\|OMP\|^$" $O/b2.err | tail -3
B2R_BENCH_GRAPH=0 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29547 bench.py --gpus 2 --steps 20 --warmup 5 --no-collectives 2> $O/b2e.err | python -c "
import json,sys; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('N=2 eager', round(d['value']), d['ms_per_step'])"
