#!/usr/bin/env python3
"""Which kernels did a change touch? Compares the SASS of two builds of libb2r.so function by function.

    git worktree add build/wt <validated-commit> && make -C build/wt -j lib
    scripts/sass_compare.py build/wt/cuda_mpi_reductions_b200/lib/libb2r.so cuda_mpi_reductions_b200/lib/libb2r.so [function]

Instruction streams are compared after normalising what legitimately moves when a kernel-parameter struct grows
(constant-bank offsets of parameters, `sizeof` multipliers given with --size-imm, branch targets). Used at the end of
round 1, when kernels were added without a GPU to re-validate on: all 240 kernels of the last GPU-tested commit came out
identical, so the new code cannot have changed what the tests had passed on. With a function name, prints the first
differing instructions of that function."""
import argparse
import collections
import re
import subprocess


def dump(lib, size_imms):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
    funcs, cur = collections.OrderedDict(), None
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = re.sub(r"_GLOBAL__N__[0-9a-f]+_", "_GLOBAL__N__", m.group(1))  # anonymous-namespace names carry a path hash
            funcs[cur] = []
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", line) if cur else None
        if not m:
            continue
        ins = m.group(1)
        ins = re.sub(r"c\[0x0\]\[0x[0-9a-f]+\]", "c[0x0][P]", ins)
        if ins.startswith(("BRA", "CALL", "BSSY", "BSYNC", "WARPSYNC")) or " `(" in ins:
            ins = re.sub(r"0x[0-9a-f]+", "IMM", ins)
        if size_imms:
            ins = re.sub(r"(U?IMAD \S+ \S+ )0x(%s)," % "|".join(size_imms), r"\1SZ,", ins)
        funcs[cur].append(ins)
    return funcs


def main():
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument("old")
    ap.add_argument("new")
    ap.add_argument("function", nargs="?")
    ap.add_argument("--size-imm", default="", help="comma-separated hex immediates (no 0x) to treat as equal struct sizes, e.g. d8,e8")
    args = ap.parse_args()
    imms = [x for x in args.size_imm.split(",") if x]
    a, b = dump(args.old, imms), dump(args.new, imms)
    same = diff = 0
    for f, ia in a.items():
        if f not in b:
            print("only in old:", f)
        elif ia == b[f]:
            same += 1
        else:
            diff += 1
            ha, hb = (collections.Counter(x.split()[0] for x in v) for v in (ia, b[f]))
            print(f"DIFF {f}: {len(ia)} -> {len(b[f])} instructions, opcode histogram {'equal' if ha == hb else 'differs'}")
    print(f"identical {same}, different {diff}, only in new {len([f for f in b if f not in a])}")
    if args.function and args.function in a and args.function in b:
        shown = 0
        for x, y in zip(a[args.function], b[args.function]):
            if x != y and shown < 10:
                print("  old", x)
                print("  new", y)
                shown += 1
    return 1 if diff else 0


if __name__ == "__main__":
    raise SystemExit(main())
