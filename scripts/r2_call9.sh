#!/bin/bash
# final 1-GPU check of HEAD: GPU test suite, smoke(), both bench arms
O=gpurun_out/r2c9; mkdir -p $O
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 25 > $O/pytest.txt; tail -n 6 $O/pytest.txt
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.txt 2>&1; echo "smoke rc=$?"; tail -n 2 $O/smoke.txt
timeout 300 python bench.py --impl reference --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$?"
timeout 300 python bench.py --steps 20 --warmup 5 > $O/bench.json 2> $O/bench.err; echo "bench rc=$?"
tail -c 1500 $O/bench_ref.json; echo; head -c 2500 $O/bench.json
