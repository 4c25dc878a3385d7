#!/usr/bin/env python
"""posted_table.py plan.csv nccl.csv — per (dtype, op, size): previous choice (Q.base), posted, NCCL, in us per call, and ratios."""
import csv
import sys
from collections import defaultdict

rows = defaultdict(dict)
for path in sys.argv[1:]:
    for r in csv.reader(open(path)):
        if len(r) < 10 or r[1] != "reduce":
            continue
        key = (r[2], r[3], int(r[6]))
        if r[0] == "nccl":
            rows[key]["nccl"] = float(r[7])
        else:
            label = r[10] if len(r) > 10 else r[4]
            rows[key][label] = float(r[7])
            if label == "Q.base":
                rows[key]["base_algo"] = r[4]
print(f"{'dtype':7s}{'op':5s}{'bytes':>10s} {'base':>9s} {'(algo)':>8s} {'posted':>9s} {'nccl':>9s} {'posted/base':>12s} {'nccl/posted':>12s} {'nccl/base':>10s}")
for (dt, op, b), v in sorted(rows.items()):
    base, post, nccl = v.get("Q.base"), v.get("Q.posted"), v.get("nccl")
    f = lambda x: f"{x:9.2f}" if x else "        -"
    g = lambda a, c: f"{a / c:12.2f}" if a and c else "           -"
    print(f"{dt:7s}{op:5s}{b:10d} {f(base)} {v.get('base_algo', '-'):>8s} {f(post)} {f(nccl)} {g(post, base)} {g(nccl, post)} {g(nccl, base)[2:]}")
