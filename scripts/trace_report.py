#!/usr/bin/env python
"""Phase breakdown of one reduction launch from the in-kernel %globaltimer stamps (B2R trace).
Usage: python scripts/trace_report.py [--n 16777216] [--dtype int32] [--variant ldg128] [--threads 512] [--ctas 592]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import cuda_mpi_reductions_b200 as b2r  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 24)
ap.add_argument("--dtype", default="int32")
ap.add_argument("--variant", default="auto")
ap.add_argument("--threads", type=int, default=0)
ap.add_argument("--ctas", type=int, default=0)
ap.add_argument("--cluster", type=int, default=0)
ap.add_argument("--stages", type=int, default=0)
ap.add_argument("--pdl", type=int, default=0)
a = ap.parse_args()
dtype = getattr(torch, a.dtype)
cfg = {k: v for k, v in dict(variant=a.variant, threads=a.threads, ctas=a.ctas, cluster=a.cluster, stages=a.stages, pdl=a.pdl).items() if v not in (0, "auto")}
ws = b2r.Workspace()
xs = [b2r.fill_device(a.n, dtype, seed=i) for i in range(10)]
for x in xs:
    b2r.reduce(x, "sum", workspace=ws, cfg=cfg)
torch.cuda.synchronize()
ws.trace(True)
ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
ev0.record()
b2r.reduce(xs[0], "sum", workspace=ws, cfg=cfg)
ev1.record()
torch.cuda.synchronize()
t = b2r.trace_last(ws).double()
t0 = t[:, 0].min()
t = (t - t0) / 1e3
print(f"plan: {b2r.plan(a.n, dtype, cfg=cfg)}  event time {ev0.elapsed_time(ev1)*1e3:.2f} us")
names = ["start", "streamed", "combined", "exit"]
for i, nm in enumerate(names):
    c = t[:, i]
    print(f"{nm:9s} min {c.min():7.2f}  p50 {c.median():7.2f}  p95 {c.quantile(0.95):7.2f}  max {c.max():7.2f} us")
dur = t[:, 1] - t[:, 0]
print(f"stream duration per CTA: min {dur.min():.2f} p50 {dur.median():.2f} max {dur.max():.2f} us")
