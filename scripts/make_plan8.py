#!/usr/bin/env python
"""Writes the measurement plan of the 8-GPU tuning call (bench_collective --plan): every candidate algorithm and launch shape
for reduce-to-root and allreduce, so that ONE process (one communicator set-up) decides b2r_select_algo's crossovers."""
import sys

K, M, G = 1 << 10, 1 << 20, 1 << 30
S_ALL = [1 * K, 4 * K, 16 * K, 64 * K, 256 * K, 1 * M, 4 * M, 16 * M, 64 * M, 256 * M, 1 * G]
rows = []


def add(label, coll, dtype, op, algo, nbytes, **knobs):
    rows.append(f"{label} {coll} {dtype} {op} {algo} {nbytes} " + " ".join(f"{k}={v}" for k, v in knobs.items()))


# ---- A. reduce fp32 SUM: every algorithm over the size range
for s in S_ALL:
    add("A.auto", "reduce", "FLOAT", "SUM", "auto", s)
    add("A.nvls", "reduce", "FLOAT", "SUM", "nvls", s)
    if s <= 256 * K:
        add("A.ll", "reduce", "FLOAT", "SUM", "ll", s)
    if s <= 4 * M:
        add("A.oneshot", "reduce", "FLOAT", "SUM", "oneshot", s)
    if s >= 64 * K:
        add("A.twoshot", "reduce", "FLOAT", "SUM", "twoshot", s)
    if s >= 256 * K:
        add("A.ring", "reduce", "FLOAT", "SUM", "ring", s)
        add("A.ring2", "reduce", "FLOAT", "SUM", "ring", s, B2R_RING_V2=1)
    if s >= 1 * M and s <= 64 * M:
        add("A.push", "reduce", "FLOAT", "SUM", "push", s)
    if s <= 16 * M:
        add("A.tree", "reduce", "FLOAT", "SUM", "tree", s)
# NVLS reduce: uneven split (root pulls share/64 of the message itself) and launch shapes
for s in (4 * M, 16 * M, 64 * M, 256 * M, 1 * G):
    for share in (16, 24, 32, 48, 64):
        add(f"A.nvls.share{share}", "reduce", "FLOAT", "SUM", "nvls", s, B2R_NVLS_ROOT_SHARE=share)
for s in (256 * M,):
    for ctas in (32, 64, 128):
        for thr in (256, 512):
            for share in (0, 32):
                add(f"A.nvls.shape", "reduce", "FLOAT", "SUM", "nvls", s, B2R_COLL_CTAS=ctas, B2R_COLL_THREADS=thr, B2R_NVLS_ROOT_SHARE=share)
# ---- B. ring launch shapes (both versions) at two sizes
for s in (16 * M, 256 * M):
    for v2 in (0, 1):
        for ctas in (16, 32, 64, 128):
            for subs in (1, 2, 4, 8):
                for thr in (256, 512):
                    add(f"B.ring{2 if v2 else 1}", "reduce", "FLOAT", "SUM", "ring", s, B2R_RING_V2=v2, B2R_RING_CTAS=ctas, B2R_RING_SUBBLOCKS=subs, B2R_COLL_THREADS=thr)
# ---- C. reduce int32 SUM / MIN (no 16-byte multimem form) and fp32 MIN (no multimem form at all)
for dtype, op in (("INT", "SUM"), ("INT", "MIN"), ("FLOAT", "MIN")):
    for s in S_ALL:
        add("C.auto", "reduce", dtype, op, "auto", s)
        if dtype == "INT":
            add("C.nvls", "reduce", dtype, op, "nvls", s)
        if s <= 256 * K:
            add("C.ll", "reduce", dtype, op, "ll", s)
        if s <= 4 * M:
            add("C.oneshot", "reduce", dtype, op, "oneshot", s)
        if s >= 64 * K:
            add("C.twoshot", "reduce", dtype, op, "twoshot", s)
            add("C.twoshot.u4", "reduce", dtype, op, "twoshot", s, B2R_COLL_UNROLL=4)
        if s >= 256 * K:
            add("C.ring", "reduce", dtype, op, "ring", s)
            add("C.ring2", "reduce", dtype, op, "ring", s, B2R_RING_V2=1)
# ---- D. allreduce
for s in S_ALL:
    add("D.auto", "allreduce", "FLOAT", "SUM", "auto", s)
    if s <= 256 * K:
        add("D.ll", "allreduce", "FLOAT", "SUM", "ll", s)
        add("D.oneshot", "allreduce", "FLOAT", "SUM", "oneshot", s)
    if s >= 64 * K:
        add("D.twoshot", "allreduce", "FLOAT", "SUM", "twoshot", s)
for ctas in (16, 32, 64, 128):
    for thr in (128, 256, 512):
        for unr in (4, 8):
            for s in (256 * M, 1 * G):
                add("D.nvls.shape", "allreduce", "FLOAT", "SUM", "nvls", s, B2R_COLL_CTAS=ctas, B2R_COLL_THREADS=thr, B2R_COLL_UNROLL=unr)
for dtype, op in (("INT", "SUM"), ("INT", "MIN"), ("FLOAT", "MIN")):
    for s in S_ALL:
        add("D.auto", "allreduce", dtype, op, "auto", s)
        if dtype == "INT":
            add("D.nvls", "allreduce", dtype, op, "nvls", s)
        if s <= 256 * K:
            add("D.ll", "allreduce", dtype, op, "ll", s)
            add("D.oneshot", "allreduce", dtype, op, "oneshot", s)
        if s >= 64 * K:
            add("D.twoshot", "allreduce", dtype, op, "twoshot", s)
        if s >= 1 * M:
            add("D.ringall", "allreduce", dtype, op, "ring", s, B2R_EXPERIMENTAL=1)
for ctas in (32, 64, 128):
    for subs in (1, 2, 4, 8):
        add("D.ringall.shape", "allreduce", "INT", "SUM", "ring", 256 * M, B2R_EXPERIMENTAL=1, B2R_RING_CTAS=ctas, B2R_RING_SUBBLOCKS=subs)
# ---- E. the other NVLS types, AUTO
for dtype in ("DOUBLE", "BF16"):
    for coll in ("reduce", "allreduce"):
        for s in S_ALL:
            add("E.auto", coll, dtype, "SUM", "auto", s)
            if s <= 256 * K:
                add("E.ll", coll, dtype, "SUM", "ll", s)
        for s in (16 * M, 256 * M):
            add("E.ring2", "reduce", dtype, "SUM", "ring", s, B2R_RING_V2=1)
out = sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/plan8.txt"
with open(out, "w") as f:
    f.write("# generated by scripts/make_plan8.py\n" + "\n".join(rows) + "\n")
print(len(rows), "rows ->", out)
