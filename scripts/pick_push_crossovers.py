#!/usr/bin/env python
"""Reads the plan CSV of the push measurement and prints shell exports for the three push crossovers: the smallest size from
which push beats AUTO at every larger measured size, per class (0 = never)."""
import collections
import csv
import sys

rows = collections.defaultdict(dict)
for r in csv.reader(open(sys.argv[1])):
    if len(r) > 10 and r[0] == "b2r" and r[10] in ("P.auto", "P.push"):
        rows[(r[1], r[2], r[3])][(int(r[6]), r[10])] = float(r[7])
NVLS = {("FLOAT", "SUM"), ("INT", "SUM"), ("DOUBLE", "SUM"), ("BF16", "SUM"), ("INT", "MIN")}


def crossover(keys):
    sizes = sorted({s for k in keys for (s, _) in rows.get(k, {})})
    best = 0
    for s in reversed(sizes):
        ok = all(rows[k].get((s, "P.push"), 1e30) < 0.97 * rows[k].get((s, "P.auto"), 0) for k in keys if k in rows)
        if not ok:
            break
        best = s
    return best >> 20


red_nvls = crossover([("reduce", d, o) for d, o in NVLS])
red_other = crossover([("reduce", "FLOAT", "MIN"), ("reduce", "DOUBLE", "MAX")])
all_other = crossover([("allreduce", "FLOAT", "MIN"), ("allreduce", "DOUBLE", "MAX")])
print(f"export B2R_AUTO_PUSH_REDUCE_NVLS_MB={red_nvls} B2R_AUTO_PUSH_REDUCE_MB={red_other} B2R_AUTO_PUSH_ALLREDUCE_MB={all_other}")
