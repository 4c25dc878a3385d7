#!/bin/bash
O=gpurun_out/r2c4; mkdir -p $O
timeout 900 python -m pytest tests/test_reduce_local_gpu.py tests/test_zz_no_alloc.py tests/test_package_api.py tests/test_collectives_loopback_gpu.py -m gpu -x -q 2>&1 | tail -8 > $O/pytest.txt; tail -3 $O/pytest.txt
for pf in 0 1 2 4 8; do echo "== prefetch lines $pf"; B2R_PREFETCH_LINES=$pf timeout 200 bin/local_sweep --pdl=1 --dyn --iters=200 2>&1 | head -6; done > $O/prefetch_sweep_64m.txt; grep -A4 "==" $O/prefetch_sweep_64m.txt | grep -v "^#\|variant"
for pf in 0 2; do echo "== prefetch lines $pf"; B2R_PREFETCH_LINES=$pf timeout 200 bin/local_sweep --pdl=1 --iters=200 2>&1 | head -10; done > $O/prefetch_geom_64m.txt; grep -A6 "==" $O/prefetch_geom_64m.txt | grep -v "^#\|variant"
echo "== double"; B2R_PREFETCH_LINES=2 timeout 200 bin/local_sweep --pdl=1 --dyn --type=double --n=8388608 --iters=200 2>&1 | head -5 | tee $O/f64.txt
echo "== 128m"; B2R_PREFETCH_LINES=2 timeout 200 bin/local_sweep --pdl=1 --dyn --n=33554432 --iters=100 2>&1 | head -5 | tee $O/128m.txt
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench20.json 2> $O/bench20.err; python -c "
import json; d=json.load(open('$O/bench20.json')); print(d['value'], d['value_plain_stream_order'], d['clocks'])"
timeout 100 baseline/_ref/bin/ref_e2e --steps=20; timeout 100 baseline/_ref/bin/ref_e2e --steps=20 --pageable
