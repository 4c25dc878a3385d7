#!/usr/bin/env python
"""Plan for the posted reduce: algo=posted against what AUTO picked before it existed (posted switched off), 512 KiB - 64 MiB,
for every dtype class; plus the AUTO rows as shipped. Usage: make_plan_posted.py > scripts/plans/plan_posted.txt"""
K, M = 1 << 10, 1 << 20
OFF = " ".join(f"B2R_AUTO_POSTED_{k}=0" for k in ("MAX_BYTES", "KEY_MAX_BYTES_8", "KEY_MAX_BYTES_4"))
for dtype, op in (("FLOAT", "SUM"), ("INT", "SUM"), ("DOUBLE", "SUM"), ("BF16", "SUM"), ("FLOAT", "MIN"), ("DOUBLE", "MAX"), ("INT", "MAX")):
    for s in (512 * K, 1 * M, 2 * M, 4 * M, 8 * M, 16 * M, 32 * M, 64 * M):
        print(f"Q.base reduce {dtype} {op} auto {s} {OFF}")
        print(f"Q.posted reduce {dtype} {op} posted {s}")
