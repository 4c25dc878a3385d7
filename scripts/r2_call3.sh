#!/bin/bash
O=gpurun_out/r2c3; mkdir -p $O
bin/launch_gap > $O/launch_gap.txt 2>&1; cat $O/launch_gap.txt
timeout 300 bin/local_sweep --pdl=1 --iters=200 > $O/sweep_pdl1_64m.txt 2>&1; head -16 $O/sweep_pdl1_64m.txt
timeout 300 bin/local_sweep --pdl=1 --iters=100 --n=33554432 > $O/sweep_pdl1_128m.txt 2>&1; head -8 $O/sweep_pdl1_128m.txt
B2R_DYN_PCT=40 timeout 300 bin/local_sweep --pdl=1 --dyn --iters=200 > $O/sweep_pdl1_dyn_64m.txt 2>&1; head -8 $O/sweep_pdl1_dyn_64m.txt
