"""b200reduce — Blackwell-native (sm_100a) sum/min/max reductions and MPI-style reduce/allreduce over NVLink.

Python face of the C/CUDA core in `include/b2r.h` (`lib/libb2r.so`). Capabilities mirror the reference
szabodabo/CUDA-MPI-Reductions: the single-GPU reduction sample (cuda/C/src/reduction) and the MPI_Reduce
benchmark (mpi/reduce.c); see SURVEY.md for the full map.
"""
from . import _lib
from ._lib import B2RError, build, refresh_env, PDL_OFF, PDL_SAFE, PDL_INPUT_READY
from .ops.local import Workspace, reduce, reduce_from_host, dot, plan, trace_last, launch_config
from .ops.oracle import reduce_oracle, combine_oracle, fill_host, fill_device

__all__ = [
    "B2RError", "build", "refresh_env", "Workspace", "reduce", "reduce_from_host", "dot", "plan", "trace_last", "launch_config",
    "reduce_oracle", "combine_oracle", "fill_host", "fill_device",
]
__version__ = "0.1.0"
