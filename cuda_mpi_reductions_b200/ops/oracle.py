"""CPU oracles and synthetic-input generation, backed by the C core (src/core/oracle.cpp, src/core/rng.h).

Reference counterparts: sumreduceCPU / minreduceCPU / maxreduceCPU (reduction.cpp:214-249) and the input fills
(reduction.cpp:698-705, mpi/reduce.c:51-57). Semantics are spelled out in include/b2r.h.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib

_TORCH2B2R = {torch.int32: _lib.INT32, torch.float32: _lib.FLOAT32, torch.float64: _lib.FLOAT64,
              torch.bfloat16: _lib.BFLOAT16}
_B2R2TORCH = {v: k for k, v in _TORCH2B2R.items()}
FILL_MODES = {"sdk": 0, "full": 1, "signed": 2}


def b2r_dtype(t: torch.dtype) -> int:
    try:
        return _TORCH2B2R[t]
    except KeyError:
        raise TypeError(f"unsupported dtype {t}; supported: int32, float32, float64, bfloat16") from None


def op_code(op) -> int:
    if isinstance(op, int):
        return op
    try:
        return _lib.OPS[op]
    except KeyError:
        raise ValueError(f"unknown op {op!r}; expected sum|min|max") from None


def reduce_oracle(x: torch.Tensor, op="sum") -> torch.Tensor:
    """Scalar CPU oracle for `reduce`: int32 wrap-around sum (bit-exact), compensated fp sums, NaN-propagating
    min/max. Returns a 0-dim CPU tensor of x.dtype."""
    x = x.detach().cpu().contiguous().reshape(-1)
    out = torch.empty((), dtype=x.dtype)
    st = _lib.LIB.b2r_reduce_host(C.c_void_p(x.data_ptr() if x.numel() else 0), C.c_void_p(out.data_ptr()), x.numel(),
                                  b2r_dtype(x.dtype), op_code(op))
    _lib.check(st, "b2r_reduce_host")
    return out


def combine_oracle(acc: torch.Tensor, x: torch.Tensor, op="sum") -> torch.Tensor:
    """In-place element-wise acc = op(acc, x) on CPU tensors (oracle for reduce/allreduce)."""
    assert acc.device.type == "cpu" and x.device.type == "cpu" and acc.dtype == x.dtype and acc.numel() == x.numel()
    assert acc.is_contiguous()
    x = x.contiguous()
    st = _lib.LIB.b2r_combine_host(C.c_void_p(acc.data_ptr()), C.c_void_p(x.data_ptr()), acc.numel(), b2r_dtype(acc.dtype),
                                   op_code(op))
    _lib.check(st, "b2r_combine_host")
    return acc


def fill_host(n: int, dtype: torch.dtype, seed: int = 0, first_index: int = 0, mode: str = "sdk",
              pin_memory: bool = False) -> torch.Tensor:
    """Synthetic input on the CPU: element i = f(seed, first_index + i); bit-identical to fill_device."""
    t = torch.empty(n, dtype=dtype, pin_memory=pin_memory)
    st = _lib.LIB.b2r_fill_host(C.c_void_p(t.data_ptr() if n else 0), n, b2r_dtype(dtype), seed, first_index, FILL_MODES[mode])
    _lib.check(st, "b2r_fill_host")
    return t


def fill_device(n: int, dtype: torch.dtype, device="cuda", seed: int = 0, first_index: int = 0, mode: str = "sdk",
                out: torch.Tensor | None = None) -> torch.Tensor:
    """Synthetic input generated on the GPU by the library's fill kernel."""
    t = out if out is not None else torch.empty(n, dtype=dtype, device=device)
    assert t.is_cuda and t.is_contiguous() and t.numel() == n and t.dtype == dtype
    with torch.cuda.device(t.device):
        stream = torch.cuda.current_stream().cuda_stream
        st = _lib.LIB.b2r_fill_device(C.c_void_p(t.data_ptr() if n else 0), n, b2r_dtype(dtype), seed, first_index,
                                      FILL_MODES[mode], C.c_void_p(stream))
    _lib.check(st, "b2r_fill_device")
    return t
