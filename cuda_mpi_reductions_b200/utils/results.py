"""Result pipeline: the reference's mpi/getAvgs.sh (:3-14) and mpi/makePlots.gp (:1-39) restated.

  aggregate  collected rows "DTYPE OP NODES GB/sec" (what bench_collective / reduce.c print) -> results/<DTYPE>_<OP>.txt
             with the reference's layout: first line blank, then "DTYPE OP NODES AVG" per distinct node count, 5 decimals
             (bc scale=5 truncates; so do we), node counts in ascending order.
  plot       writes a gnuplot script with the same content as the reference's figures: one figure per datatype, the per-op curves over the
             number of ranks plus three horizontal lines for the single-GPU throughput (the reference hard-codes them from
             CUdata.txt; here they are read from a bench_reduce JSONL or passed explicitly). gnuplot itself is not on the
             image: the script + data are the deliverable.

  figures    the same two figures rendered directly as SVG by a small pure-Python writer (int.svg, double.svg), for machines like
             this image that have neither gnuplot nor matplotlib.

CLI:  python -m cuda_mpi_reductions_b200.utils.results figures --results results --single-gpu results/single_gpu_plain.jsonl
      python -m cuda_mpi_reductions_b200.utils.results aggregate --collected results/collected.txt --out results
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


def read_result_file(path: str):
    """[(ranks, value)] from a results/<DTYPE>_<OP>.txt file."""
    pts = []
    try:
        with open(path) as f:
            for _dt, _op, nodes, val in parse_rows(f):
                pts.append((nodes, float(val)))
    except OSError:
        pass
    return sorted(pts)


def write_svg(series: dict, out_path: str, title: str, xlabel: str = "GPUs (ranks)", ylabel: str = "bandwidth of the whole problem [GiB/s]",
              hlines: dict | None = None, logy: bool = True, width: int = 900, height: int = 560, xlabels: dict | None = None) -> str:
    """The reference's figures (mpi/makePlots.gp:21-39 -> int.eps / double.eps) rendered without gnuplot or matplotlib, neither
    of which is on the image: a self-contained SVG line chart. `series` = {label: [(x, y), ...]} drawn as lines with markers
    over a categorical x axis (the distinct x values, like the reference's 64 / 256 / 1024 ranks); `hlines` = {label: y} drawn
    as dashed horizontal lines (the single-GPU throughputs the reference pasted in from CUdata.txt)."""
    import math

    hlines = hlines or {}
    xs = sorted({x for pts in series.values() for x, _ in pts})
    ys = [y for pts in series.values() for _, y in pts if y > 0] + [y for y in hlines.values() if y > 0]
    if not xs or not ys:
        raise ValueError("nothing to plot")
    palette = ["#2ca02c", "#1f77b4", "#d62728", "#9467bd", "#ff7f0e", "#8c564b", "#17becf", "#7f7f7f", "#bcbd22", "#e377c2"]
    ml, mr, mt, mb = 80, 250, 50, 60
    pw, ph = width - ml - mr, height - mt - mb
    lo, hi = min(ys), max(ys)
    if logy:
        lo, hi = 10 ** math.floor(math.log10(lo)), 10 ** math.ceil(math.log10(hi))
        if lo == hi:
            hi = lo * 10
        fy = lambda v: mt + ph * (1 - (math.log10(max(v, lo)) - math.log10(lo)) / (math.log10(hi) - math.log10(lo)))  # noqa: E731
        ticks = [10 ** e for e in range(int(math.log10(lo)), int(math.log10(hi)) + 1)]
    else:
        lo, hi = 0.0, hi * 1.08
        fy = lambda v: mt + ph * (1 - (v - lo) / (hi - lo))  # noqa: E731
        ticks = [hi * k / 5 for k in range(6)]
    fx = lambda x: ml + pw * ((xs.index(x) + 0.5) / len(xs))  # noqa: E731
    esc = lambda t: str(t).replace("&", "&amp;").replace("<", "&lt;").replace(">", "&gt;")  # noqa: E731
    o = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{width}" height="{height}" viewBox="0 0 {width} {height}" font-family="Helvetica,Arial,sans-serif" font-size="13">',
         f'<rect width="{width}" height="{height}" fill="white"/>',
         f'<text x="{ml + pw / 2}" y="24" text-anchor="middle" font-size="15">{esc(title)}</text>',
         f'<rect x="{ml}" y="{mt}" width="{pw}" height="{ph}" fill="none" stroke="#333"/>']
    for t in ticks:
        y = fy(t)
        o.append(f'<line x1="{ml}" y1="{y:.1f}" x2="{ml + pw}" y2="{y:.1f}" stroke="#ddd"/>')
        o.append(f'<text x="{ml - 8}" y="{y + 4:.1f}" text-anchor="end">{t:g}</text>')
    for x in xs:
        o.append(f'<text x="{fx(x):.1f}" y="{mt + ph + 20}" text-anchor="middle">{esc((xlabels or {}).get(x, x))}</text>')
    o.append(f'<text x="{ml + pw / 2}" y="{height - 14}" text-anchor="middle">{esc(xlabel)}</text>')
    o.append(f'<text transform="translate(18,{mt + ph / 2}) rotate(-90)" text-anchor="middle">{esc(ylabel)}</text>')
    ly = mt + 10
    for k, (label, pts) in enumerate(series.items()):
        c = palette[k % len(palette)]
        pts = [(x, y) for x, y in sorted(pts) if y > 0]
        if not pts:
            continue
        path = " ".join(f"{'M' if i == 0 else 'L'}{fx(x):.1f},{fy(y):.1f}" for i, (x, y) in enumerate(pts))
        o.append(f'<path d="{path}" fill="none" stroke="{c}" stroke-width="2"/>')
        for x, y in pts:
            o.append(f'<circle cx="{fx(x):.1f}" cy="{fy(y):.1f}" r="4" fill="{c}"><title>{esc(label)}: {x} -> {y:.3f}</title></circle>')
        o.append(f'<line x1="{ml + pw + 12}" y1="{ly}" x2="{ml + pw + 40}" y2="{ly}" stroke="{c}" stroke-width="2"/><circle cx="{ml + pw + 26}" cy="{ly}" r="4" fill="{c}"/>')
        o.append(f'<text x="{ml + pw + 46}" y="{ly + 4}">{esc(label)}</text>')
        ly += 20
    for k, (label, y) in enumerate(hlines.items()):
        c = palette[(k + len(series)) % len(palette)]
        o.append(f'<line x1="{ml}" y1="{fy(y):.1f}" x2="{ml + pw}" y2="{fy(y):.1f}" stroke="{c}" stroke-width="2" stroke-dasharray="7,5"/>')
        o.append(f'<line x1="{ml + pw + 12}" y1="{ly}" x2="{ml + pw + 40}" y2="{ly}" stroke="{c}" stroke-width="2" stroke-dasharray="7,5"/>')
        o.append(f'<text x="{ml + pw + 46}" y="{ly + 4}">{esc(label)}</text>')
        ly += 20
    o.append("</svg>")
    with open(out_path, "w") as f:
        f.write("\n".join(o))
    return out_path


def write_figures(results_dir: str, single_gpu: dict | None = None, dtypes=("INT", "DOUBLE"), families=(("", "b2r"), ("nccl", "NCCL"), ("tree", "custom tree"))) -> list:
    """int.svg / double.svg — the reference's two figures (writeup.tex:21-29): per datatype, the element-wise reduce-to-root
    bandwidth over the number of ranks for every operation and every implementation family found under results/
    (<results>/<DT>_<OP>.txt = this library, <results>/nccl/ = NCCL, <results>/tree/ = the hand-written tree), plus the
    single-GPU scalar reduction as horizontal lines."""
    single_gpu = single_gpu or {}
    written = []
    for dt in dtypes:
        series = {}
        for sub, name in families:
            for op in OPS:
                pts = read_result_file(os.path.join(results_dir, sub, f"{dt}_{op}.txt"))
                if pts:
                    series[f"{name} reduce {op}"] = pts
        if not series:
            continue
        h = {f"one GPU {op} (scalar)": single_gpu[(dt, op)] / 1.073741824 for op in OPS if (dt, op) in single_gpu}
        written.append(write_svg(series, os.path.join(results_dir, f"{dt.lower()}.svg"),
                                 f"{dt}: reduce to root of a 2 GiB problem over NVLink vs the one-GPU scalar reduction", hlines=h))
    return written


def _human_bytes(b: int) -> str:
    for unit, sh in (("GiB", 30), ("MiB", 20), ("KiB", 10)):
        if b >= 1 << sh and b % (1 << sh) == 0:
            return f"{b >> sh} {unit}"
    return f"{b} B"


def read_sweep_csv(paths, overrides=()) -> dict:
    """Rows of `bench_collective --sweep --csv=` / `nccl_bench --sweep --csv=` (impl, coll, dtype, op, algo, gpus, bytes, us,
    algbw, busbw[, label, knobs]) -> {(gpus, coll, impl, dtype, op): {bytes: (us, busbw)}}. Later files win over earlier
    ones; `overrides` are (path, label[, max_bytes]) tuples of `--plan` CSVs whose rows with that label (up to max_bytes) replace
    the b2r row of the same (coll, dtype, op, gpus, bytes) — how re-measured rows enter a figure."""
    import csv

    out: dict = {}

    def put(r, impl):
        key = (int(r[5]), r[1], impl, r[2], r[3])
        out.setdefault(key, {})[int(r[6])] = (float(r[7]), float(r[9]))

    for path in paths:
        with open(path) as f:
            for r in csv.reader(f):
                if len(r) >= 10 and r[0] in ("b2r", "nccl"):
                    put(r, r[0])
    for ov in overrides:
        path, label = ov[0], ov[1]
        max_bytes = int(ov[2]) if len(ov) > 2 else None  # e.g. up to where AUTO picks the algorithm the plan rows forced
        with open(path) as f:
            for r in csv.reader(f):
                if len(r) >= 11 and r[0] == "b2r" and r[10] == label and (int(r[5]), r[1], "b2r", r[2], r[3]) in out:
                    if max_bytes is None or int(r[6]) <= max_bytes:
                        put(r, "b2r")
    return out


def write_sweep_figures(rows: dict, out_dir: str, prefix: str = "sweep") -> list:
    """One SVG per (GPU count, collective): bus bandwidth over message size, this library and NCCL, one pair of lines per
    (dtype, op). The figure BASELINE config 5 asks for; the reference has no counterpart (its sweep is over ranks only)."""
    os.makedirs(out_dir, exist_ok=True)
    written = []
    for gpus, coll in sorted({(k[0], k[1]) for k in rows}):
        series = {}
        for (g, c, impl, dt, op), pts in sorted(rows.items()):
            if g != gpus or c != coll:
                continue
            name = ("b2r " if impl == "b2r" else "NCCL ") + f"{dt} {op}"
            series[name] = [(b, bw) for b, (_, bw) in sorted(pts.items())]
        sizes = sorted({b for pts in series.values() for b, _ in pts})
        path = os.path.join(out_dir, f"{prefix}_{coll}_{gpus}gpu.svg")
        written.append(write_svg(series, path, f"{coll} on {gpus} GPUs: bus bandwidth over message size (b2r AUTO vs NCCL)",
                                 xlabel="message size per rank", ylabel="bus bandwidth [GB/s]", width=1200, height=620,
                                 xlabels={b: _human_bytes(b) for b in sizes}))
    return written


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
    g = sub.add_parser("figures", help="int.svg / double.svg rendered in pure Python (no gnuplot needed)")
    g.add_argument("--results", default="results")
    g.add_argument("--single-gpu", default=None)
    w = sub.add_parser("sweeps", help="bus bandwidth over message size, b2r vs NCCL, one SVG per GPU count and collective")
    w.add_argument("csv", nargs="+", help="sweep CSVs of bench_collective / nccl_bench")
    w.add_argument("--override", action="append", default=[], metavar="PLAN_CSV:LABEL[:MAX_BYTES]", help="rows of a --plan CSV that replace re-measured b2r rows")
    w.add_argument("--out", default="results/sweeps")
    args = ap.parse_args(argv)
    if args.cmd == "aggregate":
        for (dt, op), path in aggregate(args.collected, args.out).items():
            print(f"{dt} {op} -> {path}")
    elif args.cmd == "sweeps":
        ov = [tuple(o.split(":")) for o in args.override]
        for path in write_sweep_figures(read_sweep_csv(args.csv, ov), args.out):
            print(path)
    elif args.cmd == "figures":
        sg = single_gpu_lines(args.single_gpu) if args.single_gpu else {}
        for path in write_figures(args.results, sg):
            print(path)
    else:
        sg = single_gpu_lines(args.single_gpu) if args.single_gpu else {}
        print(write_gnuplot(args.results, args.out or os.path.join(args.results, "plots.gp"), sg))


if __name__ == "__main__":
    main()
