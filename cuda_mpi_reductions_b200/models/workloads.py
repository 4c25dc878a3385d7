"""The reference's workloads as data (there are no neural models in this project: the "model" is a flat array).

  SDK_REDUCTION   the modified NVIDIA SDK `reduction` sample: one array -> one scalar, defaults from reduction.cpp:665-670.
  MPI_REDUCE      mpi/reduce.c: a fixed 2 GiB problem per datatype sharded over P ranks, element-wise reduce to root 0,
                  constants from mpi/constants.h:1-5 and the loop at reduce.c:71-99.
bench.py takes its sizes and the published baseline from here (loaded by path, so that the reference arm never imports the
package); the native harnesses are C++ and carry the same constants as flag defaults, which tests/test_results_pipeline.py
checks against this table — so the numbers quoted against BASELINE.md are always measured on the reference's own configuration.
"""
from __future__ import annotations

from dataclasses import dataclass, field


@dataclass(frozen=True)
class ReductionSampleWorkload:
    name: str = "sdk_reduction"
    elements: int = 1 << 24            # reduction.cpp:665
    dtypes: tuple = ("int", "float", "double")   # + "bf16" in this library
    ops: tuple = ("SUM", "MIN", "MAX")
    timed_iterations: int = 100        # reduction.cpp:731
    ref_threads: int = 256             # reduction.cpp:666
    ref_max_blocks: int = 64           # reduction.cpp:668
    published_gbs: dict = field(default_factory=lambda: {
        ("int", "SUM"): 90.8413, ("int", "MIN"): 90.7905, ("int", "MAX"): 90.7969,           # mpi/CUdata.txt:6-8
        ("double", "SUM"): 92.7729, ("double", "MIN"): 92.6014, ("double", "MAX"): 92.7552,  # mpi/CUdata.txt:2-4
    })

    def nbytes(self, dtype: str) -> int:
        return self.elements * {"int": 4, "float": 4, "double": 8, "bf16": 2}[dtype]


@dataclass(frozen=True)
class MpiReduceWorkload:
    name: str = "mpi_reduce"
    total_ints: int = 512 * 1024 * 1024      # mpi/constants.h:1
    total_doubles: int = 256 * 1024 * 1024   # mpi/constants.h:2
    retries: int = 5                         # mpi/constants.h:5
    ops: tuple = ("MAX", "MIN", "SUM")       # order of reduce.c:26-28
    root: int = 0
    published_ranks: tuple = (64, 256, 1024)  # Blue Gene/L VN mode, mpi/results/*.txt

    def count_per_rank(self, dtype: str, nranks: int) -> int:
        return (self.total_ints if dtype == "int" else self.total_doubles) // nranks  # reduce.c:43-44

    def sweep_bytes(self, lo: int = 1 << 10, hi: int = 1 << 32, factor: int = 4):
        """Message sizes of BASELINE.json config 5 (1 KB - 4 GB)."""
        b = lo
        while b <= hi:
            yield b
            b *= factor


SDK_REDUCTION = ReductionSampleWorkload()
MPI_REDUCE = MpiReduceWorkload()
