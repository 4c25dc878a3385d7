"""Device timing helpers: CUDA events around K launches, rotating inputs larger than L2, optional L2 flush.

Replaces the reference's gettimeofday stopwatch bracketed by cudaDeviceSynchronize
(stopwatch_linux.h:86-164, reduction.cpp:319-320,373-374) and the rdtsc()/CLOCK_RATE timer (mpi/reduce.c:75-78).
"""
from __future__ import annotations

import torch

L2_BYTES = 126 * 1024 * 1024


def rotating_copies(nbytes: int, min_total: int = 4 * L2_BYTES, max_copies: int = 64) -> int:
    """How many distinct copies of an `nbytes` input to rotate through so that no timed launch finds its data in L2."""
    if nbytes <= 0:
        return 1
    return max(2, min(max_copies, -(-min_total // nbytes)))


class L2Flusher:
    """Writes a buffer larger than L2 (alternative to rotating inputs)."""

    def __init__(self, device="cuda", nbytes: int = 2 * L2_BYTES):
        self.buf = torch.empty(nbytes, dtype=torch.uint8, device=device)

    def flush(self) -> None:
        self.buf.zero_()


def time_steps(fn, steps: int, warmup: int, stream: torch.cuda.Stream | None = None) -> float:
    """Runs fn(i) `warmup` times untimed, then `steps` times between two CUDA events. Returns total milliseconds."""
    stream = stream or torch.cuda.current_stream()
    for i in range(warmup):
        fn(i)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record(stream)
    for i in range(steps):
        fn(i)
    b.record(stream)
    torch.cuda.synchronize()
    return a.elapsed_time(b)
