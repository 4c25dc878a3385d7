"""Tensor-level ops: single-GPU reductions (local.py) and CPU oracles / data generation (oracle.py)."""
