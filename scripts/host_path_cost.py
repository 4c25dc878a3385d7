#!/usr/bin/env python3
"""Host-side cost of one launch through the Python face, measured WITHOUT a GPU (the native call is pointed at a NULL
communicator / replaced by a no-op, so only argument handling, the ctypes transition and the planner run).

    python scripts/host_path_cost.py            # prints a small table; profiles/host_path.md was written from it

Why it matters: a 64 MiB reduction is 9 us of GPU time, so a rank whose host path costs more than that per call is
launch-bound no matter how good the kernel is ("Host path")."""
import ctypes as C
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import cuda_mpi_reductions_b200 as b2r  # noqa: E402
from cuda_mpi_reductions_b200 import _lib  # noqa: E402
from cuda_mpi_reductions_b200.parallel import comm as cm  # noqa: E402

N = 200000


def per_call(f):
    t0 = time.perf_counter()
    for _ in range(N):
        f()
    return (time.perf_counter() - t0) / N * 1e6


def main():
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    chosen = _lib.LaunchCfg()
    rows = []
    rows.append(("trivial ctypes call (1 int argument)", per_call(lambda: _lib.LIB.b2r_dtype_size(0))))
    fn = _lib.LIB.b2r_reduce_scalar_sharded_ex
    rows.append(("ctypes call, 10 arguments, fails fast on a NULL communicator", per_call(lambda: fn(1, 2, 1 << 24, 0, 0, 0, 0, None, 3, C.byref(cfg)))))
    # the Python wrapper with the native call and the stream lookup stubbed out
    c = object.__new__(cm.Communicator)
    c.device, c.device_index, c._h, c.rank = torch.device("cpu"), 0, C.c_void_p(1), 0
    x, out = torch.zeros(1024, dtype=torch.int32), torch.zeros((), dtype=torch.int32)
    real_ex, real_stream = cm._SCALAR_EX, cm._raw_stream
    cm._SCALAR_EX, cm._raw_stream = (lambda *a: 0), (lambda idx: 0)
    rows.append(("Communicator.reduce_scalar_sharded, Python part only", per_call(lambda: c.reduce_scalar_sharded(x, "sum", root=0, out=out, cfg=cfg))))
    cm._SCALAR_EX, cm._raw_stream = real_ex, real_stream
    plan = _lib.LIB.b2r_plan_local
    rows.append((f"b2r_plan_local incl. re-reading 7 variables ({len(os.environ)} variables set)", per_call(lambda: plan(1 << 24, 0, 0, -1, C.byref(cfg), C.byref(chosen)))))
    refresh = _lib.LIB.b2r_refresh_env
    rows.append(("b2r_refresh_env alone (bumps a generation counter)", per_call(refresh)))
    for k in range(150):
        os.environ[f"FAKE_WORKER_VARIABLE_{k}"] = "x" * 40
    rows.append((f"b2r_plan_local with {len(os.environ)} variables set (a torchrun worker)", per_call(lambda: plan(1 << 24, 0, 0, -1, C.byref(cfg), C.byref(chosen)))))
    width = max(len(r[0]) for r in rows)
    for name, us in rows:
        print(f"{name.ljust(width)}  {us:6.2f} us")


if __name__ == "__main__":
    main()
