"""Single-GPU scalar reductions on CUDA tensors (sum / min / max over int32, fp32, fp64, bf16).

Counterpart of the reference's `sumreduce/minreduce/maxreduce<T>(size, threads, blocks, whichKernel, d_in, d_out)`
(cuda/C/src/reduction/reduction.h:15-25) — but one launch, any element count, deterministic, and the launch
geometry is chosen by the library (override with `cfg=` or the B2R_* environment variables).
"""
from __future__ import annotations

import ctypes as C
import threading

import torch

from .. import _lib
from .oracle import b2r_dtype, op_code


class Workspace:
    """Per-stream scratch (partials, ticket, trace stamps). Launches sharing a workspace must be stream-ordered."""

    def __init__(self, device: int | torch.device | None = None):
        if device is None:
            device = torch.cuda.current_device()
        self.device = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
        h = C.c_void_p()
        _lib.check(_lib.LIB.b2r_workspace_create(C.byref(h), self.device.index or 0), "b2r_workspace_create")
        self._h = h

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    def trace(self, enable: bool = True) -> None:
        _lib.check(_lib.LIB.b2r_workspace_trace_enable(self._h, int(enable)), "b2r_workspace_trace_enable")

    def close(self) -> None:
        if self._h:
            _lib.LIB.b2r_workspace_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_ws_lock = threading.Lock()
_ws_cache: dict[tuple[int, int], Workspace] = {}


def _workspace_for(device: torch.device, stream: int) -> Workspace:
    key = (device.index or 0, stream)
    with _ws_lock:
        ws = _ws_cache.get(key)
        if ws is None:
            ws = _ws_cache[key] = Workspace(device)
        return ws


def launch_config(**kw) -> _lib.LaunchCfg:
    """Builds a reusable launch configuration (variant=, threads=, ctas=, cluster=, unroll=, stages=, pdl=, dyn_pct=).
    pdl: PDL_SAFE (1) lets consecutive launches overlap their launch latency; PDL_INPUT_READY (2) additionally lets a launch
    stream its input while the previous kernel drains — only for inputs that were complete before that kernel started.
    Passing the same object to many reduce() calls avoids re-parsing a dict on every launch."""
    c = _lib.LaunchCfg()
    for k, v in kw.items():
        if k == "variant" and isinstance(v, str):
            v = _lib.VARIANTS[v]
        if not hasattr(c, k):
            raise TypeError(f"unknown launch option {k!r}")
        setattr(c, k, int(v))
    return c


def _cfg(cfg) -> _lib.LaunchCfg | None:
    if cfg is None or isinstance(cfg, _lib.LaunchCfg):
        return cfg
    return launch_config(**dict(cfg))


_raw_stream = getattr(torch._C, "_cuda_getCurrentRawStream", None)
_reduce_ex = _lib.LIB.b2r_reduce_local_ex


def _current_stream(device_index: int) -> int:
    if _raw_stream is not None:
        return _raw_stream(device_index)
    return torch.cuda.current_stream(device_index).cuda_stream


def reduce(x: torch.Tensor, op="sum", out: torch.Tensor | None = None, workspace: Workspace | None = None,
           cfg=None) -> torch.Tensor:
    """out = op over all elements of the CUDA tensor `x` (0-dim tensor of x.dtype on x's device).

    One kernel launch on the current stream of x's device, no host synchronisation, no allocation when `out` is given
    (so it can be captured in a CUDA graph). int32 sums wrap (two's complement); bf16 accumulates in fp32 and rounds
    once; min/max propagate NaN; an empty input gives the identity. The host-side path is kept thin on purpose: at
    ~10 us per 64 MiB launch the Python call itself is the bottleneck otherwise.
    """
    if not x.is_cuda:
        raise ValueError("reduce() wants a CUDA tensor; use reduce_oracle() or reduce_from_host() for host data")
    if not x.is_contiguous():
        x = x.contiguous()
    dt = b2r_dtype(x.dtype)
    dev = x.device
    if out is None:
        out = torch.empty((), dtype=x.dtype, device=dev)
    elif not (out.is_cuda and out.dtype == x.dtype and out.numel() == 1 and out.device == dev):
        raise ValueError("out must be a 1-element CUDA tensor of x's dtype on x's device")
    stream = _current_stream(dev.index)
    ws = workspace if workspace is not None else _workspace_for(dev, stream)
    c = cfg if (cfg is None or isinstance(cfg, _lib.LaunchCfg)) else _cfg(cfg)
    n = x.numel()
    st = _reduce_ex(x.data_ptr() if n else 0, out.data_ptr(), n, dt, op if isinstance(op, int) else op_code(op), ws._h, stream,
                    C.byref(c) if c is not None else None)
    if st:
        _lib.check(st, "b2r_reduce_local")
    return out


def dot(a: torch.Tensor, b: torch.Tensor, workspace: Workspace | None = None) -> torch.Tensor:
    """sum(a*b) as one fused transform-reduce kernel (vendored scalarProd analogue, scalarProd_kernel.cu:33-80)."""
    assert a.is_cuda and b.is_cuda and a.dtype == b.dtype and a.numel() == b.numel() and a.device == b.device
    a, b = a.contiguous(), b.contiguous()
    out = torch.empty((), dtype=a.dtype, device=a.device)
    with torch.cuda.device(a.device):
        stream = torch.cuda.current_stream().cuda_stream
        ws = workspace or _workspace_for(a.device, stream)
        st = _lib.LIB.b2r_dot_local(C.c_void_p(a.data_ptr()), C.c_void_p(b.data_ptr()), C.c_void_p(out.data_ptr()), a.numel(),
                                    b2r_dtype(a.dtype), ws.handle, C.c_void_p(stream))
    _lib.check(st, "b2r_dot_local")
    return out


def reduce_from_host(x: torch.Tensor, op="sum", device=None, workspace: Workspace | None = None) -> torch.Tensor:
    """End-to-end reduction of a HOST tensor (pinned memory recommended): the array is streamed to the GPU in chunks
    by the copy engine while previous chunks are being reduced; returns a 0-dim CPU tensor. Synchronous."""
    if x.is_cuda:
        raise ValueError("reduce_from_host() wants a CPU tensor")
    x = x.contiguous()
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    out = torch.empty((), dtype=x.dtype)
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        ws = workspace or _workspace_for(dev, stream)
        n = x.numel()
        st = _lib.LIB.b2r_reduce_from_host(C.c_void_p(x.data_ptr() if n else 0), C.c_void_p(out.data_ptr()), n, b2r_dtype(x.dtype),
                                           op_code(op), ws.handle, C.c_void_p(stream))
    _lib.check(st, "b2r_reduce_from_host")
    return out


def plan(n: int, dtype: torch.dtype, op="sum", device: int = -1, cfg=None) -> dict:
    """The launch geometry the dispatcher picks for (n, dtype, op). device=-1 plans for a 148-SM B200 (no GPU needed)."""
    chosen = _lib.LaunchCfg()
    c = _cfg(cfg)
    st = _lib.LIB.b2r_plan_local(n, b2r_dtype(dtype), op_code(op), device, C.byref(c) if c is not None else None, C.byref(chosen))
    _lib.check(st, "b2r_plan_local")
    names = {v: k for k, v in _lib.VARIANTS.items()}
    return {"variant": names[chosen.variant], "threads": chosen.threads, "ctas": chosen.ctas, "cluster": chosen.cluster,
            "unroll": chosen.unroll, "stages": chosen.stages, "tile_bytes": chosen.tile_bytes, "pdl": chosen.pdl, "dyn_pct": chosen.dyn_pct}


def trace_last(workspace, max_ctas: int = 4096, back: int = 0) -> torch.Tensor:
    """Phase stamps (ns, %globaltimer) of a traced launch: tensor [ctas, 4] = start, streamed, combined, exit. `back` = how many
    traced launches ago (0 = the last; the library keeps a ring of 8). `workspace` is a Workspace or a raw handle."""
    buf = (C.c_uint64 * (4 * max_ctas))()
    n = C.c_int()
    h = workspace.handle if hasattr(workspace, "handle") else workspace
    _lib.check(_lib.LIB.b2r_workspace_trace_read_at(h, back, buf, max_ctas, C.byref(n)), "b2r_workspace_trace_read_at")
    return torch.tensor(list(buf[: 4 * n.value]), dtype=torch.int64).reshape(n.value, 4)
