#!/bin/bash
# posted reduce: loopback protocol tests first, then the whole GPU suite
O=gpurun_out/r2c10; mkdir -p $O
timeout 600 python -m pytest tests/test_posted_gpu.py -x -q 2>&1 | tail -n 30 > $O/posted.txt; tail -n 15 $O/posted.txt
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -n 25 > $O/pytest.txt; tail -n 6 $O/pytest.txt
