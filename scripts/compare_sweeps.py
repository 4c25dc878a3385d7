#!/usr/bin/env python3
"""Turns sweep outputs of bin/bench_collective and bin/nccl_bench into the table that decides b2r_select_algo.

    scripts/compare_sweeps.py gpurun_out/round2_reduce_8/*.txt [--md out.md]

Both tools print rows of the form
      coll      dtype  op   algo              bytes         us  algbw GB/s  busbw GB/s
Rows are grouped by (coll, dtype, op, bytes); for every group the script prints each algorithm's time, the fastest of
ours, its ratio to NCCL (when an NCCL row for the same key is among the inputs), and — per (coll, dtype, op) — the size
ranges over which each algorithm wins, i.e. the crossover points to put into `b2r_select_algo` (or to try first through
B2R_AUTO_LL_MAX_BYTES / B2R_AUTO_RING_MIN_MB). When the same (key, algo) appears more than once the fastest run is kept.
Lines that are not sweep rows (headers, PASSED/FAILED, tuning lines) are ignored, so whole directories can be passed."""
import argparse
import collections
import sys


def parse(paths):
    rows = {}
    for path in paths:
        try:
            with open(path, errors="replace") as f:
                lines = f.readlines()
        except OSError as e:
            print(f"skipping {path}: {e}", file=sys.stderr)
            continue
        for line in lines:
            f = line.split()
            if len(f) != 8 or f[0] not in ("reduce", "allreduce"):
                continue
            try:
                nbytes, us, algbw, busbw = int(f[4]), float(f[5]), float(f[6]), float(f[7])
            except ValueError:
                continue
            key = (f[0], f[1], f[2], nbytes)
            algo = f[3]
            if algo not in rows.setdefault(key, {}) or us < rows[key][algo][0]:
                rows[key][algo] = (us, algbw, busbw)
    return rows


def human(n):
    for unit, width in (("GiB", 30), ("MiB", 20), ("KiB", 10)):
        if n >= 1 << width and n % (1 << width) == 0:
            return f"{n >> width} {unit}"
    return f"{n} B"


def summarise(rows):
    """[(coll, dtype, op, [(lo_bytes, hi_bytes, algo, worst ratio vs NCCL or None)])]"""
    groups = collections.defaultdict(list)
    for (coll, dtype, op, nbytes), algos in rows.items():
        ours = {a: v for a, v in algos.items() if a != "nccl"}
        if not ours:
            continue
        best = min(ours, key=lambda a: ours[a][0])
        ratio = algos["nccl"][0] / ours[best][0] if "nccl" in algos else None
        groups[(coll, dtype, op)].append((nbytes, best, ratio))
    out = []
    for key in sorted(groups):
        runs = []
        for nbytes, best, ratio in sorted(groups[key]):
            if runs and runs[-1][2] == best:
                lo, _, _, worst = runs[-1]
                runs[-1] = (lo, nbytes, best, worst if ratio is None else (ratio if worst is None else min(worst, ratio)))
            else:
                runs.append((nbytes, nbytes, best, ratio))
        out.append(key + (runs,))
    return out


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("files", nargs="+")
    ap.add_argument("--md", help="also write the per-size table as Markdown")
    args = ap.parse_args()
    rows = parse(args.files)
    if not rows:
        print("no sweep rows found", file=sys.stderr)
        return 1
    algos = sorted({a for v in rows.values() for a in v if a != "nccl"})
    header = ["coll", "dtype", "op", "bytes"] + [f"{a} us" for a in algos] + ["NCCL us", "best of ours", "best / NCCL"]
    table = []
    for key in sorted(rows):
        coll, dtype, op, nbytes = key
        v = rows[key]
        ours = {a: v[a] for a in v if a != "nccl"}
        best = min(ours, key=lambda a: ours[a][0]) if ours else "-"
        ratio = f"{v['nccl'][0] / ours[best][0]:.2f}x" if ours and "nccl" in v else "-"
        table.append([coll, dtype, op, human(nbytes)] + [f"{v[a][0]:.2f}" if a in v else "-" for a in algos] +
                     [f"{v['nccl'][0]:.2f}" if "nccl" in v else "-", best, ratio])
    widths = [max(len(str(r[i])) for r in [header] + table) for i in range(len(header))]
    for r in [header] + table:
        print("  ".join(str(c).ljust(w) for c, w in zip(r, widths)))
    print("\nwinners by size range (the crossovers for b2r_select_algo):")
    for coll, dtype, op, runs in summarise(rows):
        parts = []
        for lo, hi, algo, worst in runs:
            span = human(lo) if lo == hi else f"{human(lo)} .. {human(hi)}"
            parts.append(f"{span}: {algo}" + (f" (>= {worst:.2f}x NCCL)" if worst is not None else ""))
        print(f"  {coll:9s} {dtype:6s} {op:4s} " + ";  ".join(parts))
    if args.md:
        with open(args.md, "w") as f:
            f.write("| " + " | ".join(header) + " |\n|" + "---|" * len(header) + "\n")
            for r in table:
                f.write("| " + " | ".join(str(c) for c in r) + " |\n")
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
