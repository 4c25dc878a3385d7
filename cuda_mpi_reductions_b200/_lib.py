"""ctypes binding to libb2r.so (the C/CUDA core, built in-tree by `make lib`).

The core has a plain C ABI (include/b2r.h); this module declares the prototypes once and exposes the loaded
library as `LIB`. There is deliberately no pure-PyTorch fallback for GPU work: if the shared object is missing the
import of a GPU op raises, so a silent eager fallback can never masquerade as the native path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(_HERE)
LIB_PATH = os.path.join(_HERE, "lib", "libb2r.so")

# enums (mirror include/b2r.h)
OK, ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_TIMEOUT, ERR_NOMEM, ERR_SYSTEM, ERR_PENDING = range(8)
INT32, FLOAT32, FLOAT64, BFLOAT16 = range(4)
SUM, MIN, MAX = range(3)
VARIANT_AUTO, VARIANT_LDG128, VARIANT_LDG256, VARIANT_TMA = range(4)
ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_TREE, ALGO_NVLS, ALGO_PUSH, ALGO_RING, ALGO_LL, ALGO_POSTED = range(9)

OPS = {"sum": SUM, "min": MIN, "max": MAX, "SUM": SUM, "MIN": MIN, "MAX": MAX}
ALGOS = {"auto": ALGO_AUTO, "oneshot": ALGO_ONESHOT, "twoshot": ALGO_TWOSHOT, "tree": ALGO_TREE, "nvls": ALGO_NVLS, "push": ALGO_PUSH, "ring": ALGO_RING,
         "ll": ALGO_LL, "posted": ALGO_POSTED}  # "ll": small messages; "posted": mid-size reduce-to-root (stage locally, root reduces via the switch)
VARIANTS = {"auto": VARIANT_AUTO, "ldg128": VARIANT_LDG128, "ldg256": VARIANT_LDG256, "tma": VARIANT_TMA}
DTYPE_SIZE = {INT32: 4, FLOAT32: 4, FLOAT64: 8, BFLOAT16: 2}


class LaunchCfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("variant", "threads", "ctas", "cluster", "stages", "tile_bytes", "unroll", "pdl", "dyn_pct")]


PDL_OFF, PDL_SAFE, PDL_INPUT_READY = -1, 1, 2  # b2r_launch_cfg_t.pdl (include/b2r.h)


class CommConfig(C.Structure):
    _fields_ = [("heap_bytes", C.c_uint64), ("use_multicast", C.c_int), ("timeout_ms", C.c_int), ("max_ctas", C.c_int),
                ("symmetric_only", C.c_int), ("reserved", C.c_int * 3)]


class B2RError(RuntimeError):
    def __init__(self, status: int, where: str, detail: str):
        self.status = status
        super().__init__(f"{where}: status {status} ({detail})")


def build(verbose: bool = False) -> str:
    """Builds libb2r.so (and the native benches) in-tree with nvcc for sm_100a. Works without a GPU."""
    cmd = ["make", "-C", REPO_ROOT, "-j", str(min(16, os.cpu_count() or 4)), "lib", "bins"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout[-4000:] + res.stderr[-8000:])
    if res.returncode != 0:
        raise RuntimeError("building libb2r.so failed (see output above)")
    return LIB_PATH


def _declare(lib: C.CDLL) -> None:
    u64, i, vp, cp = C.c_uint64, C.c_int, C.c_void_p, C.c_char_p

    def sig(name, argtypes, restype=C.c_int):
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = restype

    sig("b2r_dtype_size", [i], C.c_size_t)
    sig("b2r_status_string", [i], cp)
    sig("b2r_last_error_string", [], cp)
    sig("b2r_version", [], i)
    sig("b2r_reduce_host", [vp, vp, u64, i, i])
    sig("b2r_combine_host", [vp, vp, u64, i, i])
    sig("b2r_fill_host", [vp, u64, i, u64, u64, i])
    sig("b2r_fill_device", [vp, u64, i, u64, u64, i, vp])
    sig("b2r_workspace_create", [C.POINTER(vp), i])
    sig("b2r_workspace_destroy", [vp])
    sig("b2r_reduce_local", [vp, vp, u64, i, i, vp, vp])
    sig("b2r_reduce_local_ex", [vp, vp, u64, i, i, vp, vp, C.POINTER(LaunchCfg)])
    sig("b2r_dot_local", [vp, vp, vp, u64, i, vp, vp])
    sig("b2r_plan_local", [u64, i, i, i, C.POINTER(LaunchCfg), C.POINTER(LaunchCfg)])
    sig("b2r_workspace_trace_enable", [vp, i])
    sig("b2r_workspace_trace_read", [vp, C.POINTER(u64), i, C.POINTER(i)])
    sig("b2r_workspace_trace_read_at", [vp, i, C.POINTER(u64), i, C.POINTER(i)])
    sig("b2r_reduce_from_host", [vp, vp, u64, i, i, vp, vp])
    for name, args in (
        ("b2r_comm_init_all", [C.POINTER(vp), i, C.POINTER(i), C.POINTER(CommConfig)]),
        ("b2r_comm_init_rank", [C.POINTER(vp), i, i, i, cp, C.POINTER(CommConfig)]),
        ("b2r_comm_init_loopback", [C.POINTER(vp), i, i, C.POINTER(CommConfig)]),
        ("b2r_guard_alloc", [u64, i, i, C.POINTER(vp)]),
        ("b2r_guard_free", [vp]),
        ("b2r_comm_destroy", [vp]),
        ("b2r_comm_rank", [vp, C.POINTER(i)]),
        ("b2r_comm_size", [vp, C.POINTER(i)]),
        ("b2r_comm_device", [vp, C.POINTER(i)]),
        ("b2r_comm_has_multicast", [vp, C.POINTER(i)]),
        ("b2r_comm_workspace", [vp, C.POINTER(vp)]),
        ("b2r_symm_alloc", [vp, u64, C.POINTER(vp)]),
        ("b2r_symm_free_all", [vp]),
        ("b2r_symm_peer_ptr", [vp, vp, i, C.POINTER(vp)]),
        ("b2r_comm_check", [vp, vp]),
        ("b2r_reduce", [vp, vp, u64, i, i, i, vp, vp]),
        ("b2r_allreduce", [vp, vp, u64, i, i, vp, vp]),
        ("b2r_reduce_ex", [vp, vp, u64, i, i, i, vp, vp, i]),
        ("b2r_allreduce_ex", [vp, vp, u64, i, i, vp, vp, i]),
        ("b2r_reduce_scalar_sharded", [vp, vp, u64, i, i, i, i, vp, vp]),
        ("b2r_reduce_scalar_sharded_ex", [vp, vp, u64, i, i, i, i, vp, vp, C.POINTER(LaunchCfg)]),
        ("b2r_reduce_scalar_sharded_from_host", [vp, vp, u64, i, i, i, i, vp, vp]),
        ("b2r_barrier", [vp, vp]),
        ("b2r_select_algo", [i, u64, i, i, i, i]),
        ("b2r_host_reduce_threads", [C.POINTER(vp), vp, u64, i, i, i, i, i]),
    ):
        if hasattr(lib, name):
            sig(name, args)
    if hasattr(lib, "b2r_ll_max_bytes"):
        sig("b2r_ll_max_bytes", [i], u64)
    if hasattr(lib, "b2r_comm_ll_max_bytes"):
        sig("b2r_comm_ll_max_bytes", [vp, C.POINTER(C.c_uint64)])
    sig("b2r_comm_posted_max_bytes", [vp, C.POINTER(C.c_uint64)])
    if hasattr(lib, "b2r_refresh_env"):
        sig("b2r_refresh_env", [], None)


def _load() -> C.CDLL:
    if not os.path.exists(LIB_PATH):
        if os.environ.get("B2R_NO_AUTOBUILD"):
            raise ImportError(f"{LIB_PATH} is missing; run `make lib` (or python -c 'import __graft_entry__ as g; g.build()')")
        build()
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    _declare(lib)
    return lib


LIB = _load()


def check(status: int, where: str) -> None:
    if status != OK:
        detail = LIB.b2r_last_error_string().decode() or LIB.b2r_status_string(status).decode()
        raise B2RError(status, where, detail)


def refresh_env() -> None:
    """Makes the library re-read its B2R_* tuning variables (it caches them: a getenv per launch costs microseconds).
    Call after changing such a variable at run time (os.environ[...] = ...)."""
    LIB.b2r_refresh_env()
