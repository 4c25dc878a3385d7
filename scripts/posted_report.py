#!/usr/bin/env python
"""posted_report.py <dir> — markdown table of the posted reduce on 8 GPUs from profiles/raw/r2_posted: what AUTO picked before
(Q.base rows of 8gpu_plan.csv), the posted reduce as shipped (8gpu_coalesced_plan.csv; bf16 rows, whose 16-byte requests the
coalescing change does not touch, from 8gpu_plan.csv) and NCCL (8gpu_nccl.csv, same call as the Q rows)."""
import csv
import sys
from collections import defaultdict

d = sys.argv[1] if len(sys.argv) > 1 else "profiles/raw/r2_posted"
rows = defaultdict(dict)


def read(path, want):
    for r in csv.reader(open(path)):
        if len(r) < 10 or r[1] != "reduce":
            continue
        key = (r[2], r[3], int(r[6]))
        label = "nccl" if r[0] == "nccl" else (r[10] if len(r) > 10 else "")
        if label in want:
            rows[key][want[label]] = float(r[7])
            if label == "Q.base":
                rows[key]["base_algo"] = r[4]


read(f"{d}/8gpu_plan.csv", {"Q.base": "base", "Q.posted": "posted_v1"})
read(f"{d}/8gpu_nccl.csv", {"nccl": "nccl"})
read(f"{d}/8gpu_coalesced_plan.csv", {"R.posted": "posted"})


def size(b):
    return f"{b >> 20} MiB" if b >= 1 << 20 else f"{b >> 10} KiB"


print("| dtype | op | bytes | before: algo | before us | posted us | NCCL us | before / NCCL | posted / NCCL |")
print("|---|---|---|---|---|---|---|---|---|")
for (dt, op, b), v in sorted(rows.items()):
    post = v.get("posted", v.get("posted_v1") if dt == "BF16" else None)
    if "base" not in v or post is None or "nccl" not in v:
        continue
    print(f"| {dt} | {op} | {size(b)} | {v['base_algo']} | {v['base']:.2f} | {post:.2f} | {v['nccl']:.2f} | {v['nccl'] / v['base']:.2f}x | **{v['nccl'] / post:.2f}x** |")
