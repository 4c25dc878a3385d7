#!/usr/bin/env python
"""Builds results/ from the raw reference-format rows of the record calls (profiles/raw/r2_{2,4,8}gpu*/refloop_*.txt):
   results/raw/{b2r,tree,nccl}_reduce_<N>gpu.txt   copies of the raw rows
   results/collected.txt, results/{nccl,tree}/collected.txt      concatenations (the reference's collected.txt)
   results/<DT>_<OP>.txt, results/nccl/..., results/tree/...     means per GPU count (mpi/getAvgs.sh), one row per GPU count
   results/int.svg, results/double.svg                           the two figures (mpi/makePlots.gp)
Usage: python scripts/collect_results.py profiles/raw/r2_2gpu profiles/raw/r2_4gpu profiles/raw/r2_8gpu_b"""
import os
import shutil
import sys

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
from cuda_mpi_reductions_b200.utils import results as R  # noqa: E402

out = os.path.join(REPO, "results")
fam = {"b2r": "", "nccl": "nccl", "tree": "tree"}
rows = {k: [] for k in fam}
for d in sys.argv[1:]:
    for k in fam:
        src = os.path.join(d, f"refloop_{k}.txt")
        if not os.path.exists(src):
            continue
        lines = [l for l in open(src) if l.split() and l.split()[0] in ("INT", "DOUBLE", "#")]
        n = next((l.split()[2] for l in lines if not l.startswith("#")), None)
        if n is None:
            continue
        os.makedirs(os.path.join(out, "raw"), exist_ok=True)
        with open(os.path.join(out, "raw", f"{k}_reduce_{n}gpu.txt"), "w") as f:
            f.writelines(lines)
        rows[k] += lines
for k, sub in fam.items():
    if not rows[k]:
        continue
    dst = os.path.join(out, sub)
    os.makedirs(dst, exist_ok=True)
    coll = os.path.join(dst, "collected.txt")
    with open(coll, "w") as f:
        f.writelines(rows[k])
    R.aggregate(coll, dst)
sg = {}
for name in ("single_gpu_plain.jsonl",):
    p = os.path.join(out, name)
    if os.path.exists(p):
        sg.update(R.single_gpu_lines(p))
print("\n".join(R.write_figures(out, sg)))
R.write_gnuplot(out, os.path.join(out, "plots.gp"), sg)
