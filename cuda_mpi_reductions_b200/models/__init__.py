"""Workload definitions (the reference has no models; its two benchmark workloads live here)."""
from .workloads import MPI_REDUCE, SDK_REDUCTION, MpiReduceWorkload, ReductionSampleWorkload

__all__ = ["MPI_REDUCE", "SDK_REDUCTION", "MpiReduceWorkload", "ReductionSampleWorkload"]
