"""HostCommunicator — the collectives on CPU tensors across PROCESSES (one host), the multi-process form of the host
loopback transport. Same API shape as Communicator (allreduce / reduce / barrier), same algorithms
(src/coll/host_algos.h: oneshot, twoshot, tree, ring; identical combination order to the device kernels), data exchanged
through one POSIX shared-memory segment (src/coll/host_shm.cpp). torch.distributed (gloo on CPU) is plumbing only: it
carries the segment's name from rank 0 to the others. This is what lets the cross-process path be tested without a GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import uuid

import torch

from .. import _lib
from ..ops.oracle import b2r_dtype, op_code

_vp, _u64, _i = C.c_void_p, C.c_uint64, C.c_int
_lib.LIB.b2r_host_comm_init.argtypes = [C.POINTER(_vp), C.c_char_p, _i, _i, _u64]
_lib.LIB.b2r_host_comm_init.restype = _i
_lib.LIB.b2r_host_comm_destroy.argtypes = [_vp]
_lib.LIB.b2r_host_comm_destroy.restype = _i
_lib.LIB.b2r_host_comm_reduce.argtypes = [_vp, _vp, _vp, _u64, _i, _i, _i, _i]
_lib.LIB.b2r_host_comm_reduce.restype = _i
_lib.LIB.b2r_host_comm_barrier.argtypes = [_vp]
_lib.LIB.b2r_host_comm_barrier.restype = _i


class HostCommunicator:
    def __init__(self, handle: C.c_void_p, rank: int, size: int):
        self._h, self.rank, self.size = handle, rank, size

    @classmethod
    def init(cls, rank: int, size: int, shm_name: str, window_bytes: int = 16 << 20) -> "HostCommunicator":
        h = _vp()
        _lib.check(_lib.LIB.b2r_host_comm_init(C.byref(h), shm_name.encode(), rank, size, window_bytes), "b2r_host_comm_init")
        return cls(h, rank, size)

    @classmethod
    def from_torch_distributed(cls, group=None, window_bytes: int = 16 << 20) -> "HostCommunicator":
        """One communicator per process of an initialised torch.distributed job (gloo is enough) on ONE host."""
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
        box = [f"/b2r-host-{uuid.uuid4().hex[:16]}-{os.getpid()}" if rank == 0 else None]
        dist.broadcast_object_list(box, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        return cls.init(rank, world, box[0], window_bytes)

    def _run(self, x: torch.Tensor, out: torch.Tensor | None, op, root: int, algo) -> None:
        if x.is_cuda:
            raise ValueError("HostCommunicator works on CPU tensors; use Communicator for CUDA tensors")
        st = _lib.LIB.b2r_host_comm_reduce(self._h, x.data_ptr() if x.numel() else 0, out.data_ptr() if out is not None else 0, x.numel(),
                                           b2r_dtype(x.dtype), op_code(op), root, _lib.ALGOS[algo] if isinstance(algo, str) else algo)
        _lib.check(st, "b2r_host_comm_reduce")

    def allreduce(self, x: torch.Tensor, op="sum", out: torch.Tensor | None = None, algo="twoshot") -> torch.Tensor:
        x = x.contiguous()
        out = torch.empty_like(x) if out is None else out
        self._run(x, out, op, -1, algo)
        return out

    def reduce(self, x: torch.Tensor, op="sum", root: int = 0, out: torch.Tensor | None = None, algo="twoshot") -> torch.Tensor | None:
        x = x.contiguous()
        if out is None and self.rank == root:
            out = torch.empty_like(x)
        self._run(x, out if self.rank == root else None, op, root, algo)
        return out if self.rank == root else None

    def barrier(self) -> None:
        _lib.check(_lib.LIB.b2r_host_comm_barrier(self._h), "b2r_host_comm_barrier")

    def close(self) -> None:
        if self._h:
            _lib.LIB.b2r_host_comm_destroy(self._h)
            self._h = None
