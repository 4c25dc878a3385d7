"""CPU-only: the Unix-socket rendezvous that carries VMM file descriptors between ranks (src/comm/bootstrap.cpp).
Replaces what mpirun/MPI_Init do for the reference (mpi/reduce.c:32-34); runs as real separate processes."""
import ctypes as C
import multiprocessing as mp
import os
import uuid

import pytest


def _worker(boot_id, rank, nranks, q):
    from cuda_mpi_reductions_b200 import _lib

    fn = _lib.LIB.b2r_bootstrap_selftest
    fn.argtypes = [C.c_char_p, C.c_int, C.c_int]
    fn.restype = C.c_int
    st = fn(boot_id.encode(), rank, nranks)
    q.put((rank, st, _lib.LIB.b2r_last_error_string().decode()))


@pytest.mark.parametrize("nranks", [1, 2, 4, 8])
def test_bootstrap_allgather_and_fd_passing(nranks):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    boot_id = f"test-{uuid.uuid4().hex}-{os.getpid()}"
    procs = [ctx.Process(target=_worker, args=(boot_id, r, nranks, q)) for r in range(nranks)]
    for p in reversed(procs):  # rank 0 (the listener) starts LAST: the others must retry until it is up
        p.start()
    results = [q.get(timeout=60) for _ in range(nranks)]
    for p in procs:
        p.join(timeout=30)
    assert sorted(r[0] for r in results) == list(range(nranks))
    for rank, st, msg in results:
        assert st == 0, f"rank {rank}: {msg}"
