#!/bin/bash
# round 2, call 2 (1 GPU): full GPU suite after the PDL / mailbox / dynamic-tail changes, smoke, bench, dynamic-tail sweep
O=gpurun_out/r2c2; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -30 > $O/pytest.txt
tail -5 $O/pytest.txt
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; tail -3 $O/smoke.txt
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench20.json 2> $O/bench20.err; tail -c 1500 $O/bench20.json; tail -3 $O/bench20.err
timeout 300 python bench.py --steps 2000 --warmup 50 > $O/bench2000.json 2> $O/bench2000.err; tail -c 600 $O/bench2000.json
timeout 300 bin/local_sweep --dyn --iters=200 > $O/dyn_sweep_64m.txt 2>&1; head -25 $O/dyn_sweep_64m.txt
timeout 300 bin/local_sweep --dyn --iters=100 --n=33554432 > $O/dyn_sweep_128m.txt 2>&1; head -12 $O/dyn_sweep_128m.txt
timeout 300 bin/local_sweep --dyn --iters=100 --type=double --n=8388608 > $O/dyn_sweep_64m_f64.txt 2>&1; head -12 $O/dyn_sweep_64m_f64.txt
timeout 200 bin/bench_reduce --method=SUM --type=int --n=16777216 --noprompt > $O/bench_reduce_int.txt 2>&1; grep -i "throughput\|kernel\|PASS\|FAIL" $O/bench_reduce_int.txt
