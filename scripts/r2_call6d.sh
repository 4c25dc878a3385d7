#!/bin/bash
O=gpurun_out/r2c6d; mkdir -p $O
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for s in 20 100 400; do for g in "" "--graph"; do $T --master-port 2950$((RANDOM%9)) scripts/xg_trace.py --steps $s $g --notrace 2>&1 | grep "us per step"; done; done
for s in 20 100; do for g in "" "--graph"; do python scripts/xg_trace.py --steps $s $g --notrace | grep "us per step"; done; done
$T --master-port 29519 scripts/xg_trace.py --steps 20 --graph 2>&1 | grep -v "^\*\|OMP\|^$" | tee $O/trace_graph_n2.txt
timeout 600 python -m pytest tests/test_guard_gpu.py -m gpu -x -q 2>&1 | tail -5
