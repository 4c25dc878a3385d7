#!/usr/bin/env python
"""Markdown table b200reduce (AUTO) vs NCCL from the CSV that bench_collective --sweep --csv and nccl_bench --sweep --csv append to.
    python scripts/sweep_table.py profiles/raw/r2_8gpu_b/sweep.csv [link_gbs=770]"""
import csv
import sys

path = sys.argv[1]
link = float(sys.argv[2]) if len(sys.argv) > 2 else 770.0
ours, theirs = {}, {}
for r in csv.reader(open(path)):
    if len(r) < 10:
        continue
    key = (r[1], r[2], r[3], int(r[5]), int(r[6]))
    (ours if r[0] == "b2r" else theirs)[key] = (r[4], float(r[7]), float(r[8]), float(r[9]))


def human(b):
    for u, s in ((1 << 30, "GiB"), (1 << 20, "MiB"), (1 << 10, "KiB")):
        if b >= u:
            return f"{b // u} {s}"
    return str(b)


print("| coll | dtype | op | GPUs | bytes | algo | b2r us | b2r busbw GB/s | NCCL us | NCCL busbw GB/s | b2r / NCCL | b2r algbw / link |")
print("|---|---|---|---|---|---|---|---|---|---|---|---|")
for key in sorted(ours):
    algo, us, alg, bus = ours[key]
    n = theirs.get(key)
    ratio = f"**{n[1] / us:.2f}x**" if n else "-"
    print(f"| {key[0]} | {key[1]} | {key[2]} | {key[3]} | {human(key[4])} | {algo} | {us:.2f} | {bus:.1f} | {n[1] if n else '-'} | {n[3] if n else '-'} | {ratio} | {alg / link:.2f} |")
