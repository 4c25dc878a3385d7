#!/bin/bash
O=gpurun_out/r2c14; mkdir -p $O
python scripts/ncu_default.py > $O/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file $O/launches.csv python scripts/ncu_default.py > $O/ncu_launches.log 2>&1
python scripts/ncu_default.py > $O/plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:reduce_ldg -s 5 -c 6 -o $O/reduce_ldg_dyn_int32_sum python scripts/ncu_default.py > $O/ncu_full.log 2>&1
tail -n 2 $O/ncu_full.log
