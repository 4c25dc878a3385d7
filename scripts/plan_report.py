#!/usr/bin/env python
"""Reads the CSVs of an 8-GPU tuning call (bench_collective --plan --csv, nccl_bench --csv) and prints, per (collective, dtype,
op, size): NCCL's time, AUTO's time, and every candidate ranked — the table b2r_select_algo's crossovers are read off.
    python scripts/plan_report.py gpurun_out/r2c7/plan.csv gpurun_out/r2c7/nccl.csv [--md out.md]"""
import collections
import csv
import sys

plan, nccl = sys.argv[1], sys.argv[2]
md = sys.argv[sys.argv.index("--md") + 1] if "--md" in sys.argv else None
N = {}
for r in csv.reader(open(nccl)):
    if len(r) >= 10 and r[0] == "nccl":
        N[(r[1], r[2], r[3], int(r[6]))] = float(r[7])
C = collections.defaultdict(list)
for r in csv.reader(open(plan)):
    if len(r) >= 10 and r[0] == "b2r":
        label = r[10] if len(r) > 10 else r[4]
        knobs = r[11].strip() if len(r) > 11 else ""
        C[(r[1], r[2], r[3], int(r[6]))].append((float(r[7]), r[4], label, knobs))


def human(b):
    for u, s in ((1 << 30, "GiB"), (1 << 20, "MiB"), (1 << 10, "KiB")):
        if b >= u:
            return f"{b // u} {s}"
    return str(b)


lines = ["| coll | dtype | op | bytes | NCCL us | AUTO (algo) us | AUTO/NCCL speed | best candidate | best us | best/NCCL speed | runners-up |", "|---|---|---|---|---|---|---|---|---|---|---|"]
for key in sorted(C, key=lambda k: (k[0], k[1], k[2], k[3])):
    cands = sorted(C[key])
    n = N.get(key)
    auto = [c for c in cands if c[2].endswith(".auto")]
    a = auto[0] if auto else None
    best = cands[0]
    ru = "; ".join(f"{c[2]}{'[' + c[3] + ']' if c[3] else ''} {c[0]:.1f}" for c in cands[1:4])
    lines.append(f"| {key[0]} | {key[1]} | {key[2]} | {human(key[3])} | {n if n else '-'} | {f'{a[0]:.1f} ({a[1]})' if a else '-'} | "
                 f"{f'{n / a[0]:.2f}x' if (a and n) else '-'} | {best[2]}{'[' + best[3] + ']' if best[3] else ''} | {best[0]:.1f} | {f'{n / best[0]:.2f}x' if n else '-'} | {ru} |")
text = "\n".join(lines)
print(text)
if md:
    open(md, "w").write(text + "\n")
