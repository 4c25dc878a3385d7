"""Utilities: device timing (timing.py), clock sampling (clocks.py), result aggregation and plotting (results.py)."""
from .clocks import ClockSampler
from .timing import L2_BYTES, L2Flusher, rotating_copies, time_steps

__all__ = ["ClockSampler", "L2_BYTES", "L2Flusher", "rotating_copies", "time_steps"]
