"""Result pipeline: the reference's mpi/getAvgs.sh (:3-14) and mpi/makePlots.gp (:1-39) restated.

  aggregate  collected rows "DTYPE OP NODES GB/sec" (what bench_collective / reduce.c print) -> results/<DTYPE>_<OP>.txt
             with the reference's layout: first line blank, then "DTYPE OP NODES AVG" per distinct node count, 5 decimals
             (bc scale=5 truncates; so do we), node counts in ascending order.
  plot       writes a gnuplot script with the same content as the reference's figures: one figure per datatype, the per-op curves over the
             number of ranks plus three horizontal lines for the single-GPU throughput (the reference hard-codes them from
             CUdata.txt; here they are read from a bench_reduce JSONL or passed explicitly). gnuplot itself is not on the
             image: the script + data are the deliverable.

CLI:  python -m cuda_mpi_reductions_b200.utils.results aggregate --collected results/collected.txt --out results
      python -m cuda_mpi_reductions_b200.utils.results plot --results results --single-gpu results/single_gpu.jsonl
"""
from __future__ import annotations

import argparse
import json
import os
from collections import defaultdict
from decimal import ROUND_DOWN, Decimal

OPS = ("SUM", "MIN", "MAX")


def parse_rows(lines):
    """Yields (dtype, op, nodes, value) from reference-format rows; comment/blank/malformed lines are skipped. Rows of
    the custom-tree benchmark carry a leading IMPL column ("TREE INT SUM 4 1.23"): it becomes part of the dtype key."""
    for line in lines:
        parts = line.split()
        if not parts or parts[0].startswith("#"):
            continue
        if len(parts) == 5:
            parts = [parts[0] + "_" + parts[1]] + parts[2:]
        if len(parts) != 4 or parts[1] not in OPS:
            continue
        try:
            yield parts[0], parts[1], int(parts[2]), Decimal(parts[3])
        except (ValueError, ArithmeticError):
            continue


def _trunc5(x: Decimal) -> str:
    """bc's `scale=5` division truncates; decimal arithmetic keeps 9.126 + 9.238 exact where binary floats would not."""
    return str(x.quantize(Decimal("0.00001"), rounding=ROUND_DOWN))


def aggregate(collected_path: str, out_dir: str) -> dict:
    groups: dict[tuple[str, str], dict[int, list[Decimal]]] = defaultdict(lambda: defaultdict(list))
    with open(collected_path) as f:
        for dt, op, nodes, val in parse_rows(f):
            groups[(dt, op)][nodes].append(val)
    os.makedirs(out_dir, exist_ok=True)
    written = {}
    for (dt, op), by_nodes in sorted(groups.items()):
        path = os.path.join(out_dir, f"{dt}_{op}.txt")
        with open(path, "w") as f:
            f.write("\n")
            for nodes in sorted(by_nodes):
                vals = by_nodes[nodes]
                f.write(f"{dt} {op} {nodes} {_trunc5(sum(vals, Decimal(0)) / len(vals))}\n")
        written[(dt, op)] = path
    return written


def single_gpu_lines(jsonl_path: str) -> dict:
    """{(dtype, op): GB/s} from bench_reduce --jsonl output (last entry per key wins)."""
    out = {}
    with open(jsonl_path) as f:
        for line in f:
            try:
                r = json.loads(line)
            except json.JSONDecodeError:
                continue
            if r.get("bench") == "reduce_local":
                name = {"int": "INT", "double": "DOUBLE", "float": "FLOAT", "bf16": "BF16"}.get(r["dtype"], r["dtype"].upper())
                out[(name, r["op"])] = r["GBps"]
    return out


def write_gnuplot(results_dir: str, out_path: str, single_gpu: dict | None = None, dtypes=("INT", "DOUBLE"), xlabel="GPUs (ranks)") -> str:
    """One figure per datatype: bandwidth of the whole problem over the number of ranks for SUM / MIN / MAX (read from
    results/<DTYPE>_<OP>.txt, columns 3 and 4 like the reference's plot) plus dashed horizontal lines for the single-GPU
    scalar reduction of the same datatype. Our own script: SVG output, log-scaled y axis, explicit GiB/s units."""
    single_gpu = single_gpu or {}
    colours = {"SUM": "#2ca02c", "MIN": "#1f77b4", "MAX": "#d62728"}
    out = [
        "# gnuplot script written by cuda_mpi_reductions_b200.utils.results.write_gnuplot",
        "# data: <results>/<DTYPE>_<OP>.txt rows 'DTYPE OP RANKS GiB/s' (first line blank)",
        "set terminal svg size 900,600 dynamic font 'Helvetica,13'",
        "set datafile missing ''",
        "set grid ytics lc rgb '#cccccc'",
        "set logscale y 10",
        f"set xlabel '{xlabel}'",
        "set ylabel 'bandwidth of the whole problem [GiB/s]'",
        "set key outside right center box",
        "",
    ]
    for dt in dtypes:
        out.append(f"set output '{dt.lower()}_reduce.svg'")
        out.append(f"set title '{dt}: element-wise reduce to root over NVLink vs one-GPU scalar reduction'")
        terms = []
        for op in OPS:
            path = os.path.join(results_dir, f"{dt}_{op}.txt")
            terms.append(f"'{path}' using 3:4 with linespoints lw 2 pt 7 lc rgb '{colours[op]}' title 'reduce {op}'")
        for op in OPS:
            v = single_gpu.get((dt, op))
            if v is not None:
                terms.append(f"{v / 1.073741824:.3f} with lines dt 2 lw 2 lc rgb '{colours[op]}' title 'one GPU {op}'")
        out.append("plot \\\n  " + ", \\\n  ".join(terms))
        out.append("")
    with open(out_path, "w") as f:
        f.write("\n".join(out))
    return out_path


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    sub = ap.add_subparsers(dest="cmd", required=True)
    a = sub.add_parser("aggregate")
    a.add_argument("--collected", required=True)
    a.add_argument("--out", default="results")
    p = sub.add_parser("plot")
    p.add_argument("--results", default="results")
    p.add_argument("--single-gpu", default=None, help="bench_reduce --jsonl file with the single-GPU numbers")
    p.add_argument("--out", default=None)
    args = ap.parse_args(argv)
    if args.cmd == "aggregate":
        for (dt, op), path in aggregate(args.collected, args.out).items():
            print(f"{dt} {op} -> {path}")
    else:
        sg = single_gpu_lines(args.single_gpu) if args.single_gpu else {}
        print(write_gnuplot(args.results, args.out or os.path.join(args.results, "plots.gp"), sg))


if __name__ == "__main__":
    main()
