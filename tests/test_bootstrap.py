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


def _silent_client(boot_id, hold_s):
    """Connects to the rendezvous socket like a rank would, then never says which rank it is."""
    import socket
    import time

    s = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
    deadline = time.time() + 15
    while True:
        try:
            s.connect(b"\0b2r-" + boot_id.encode())
            break
        except OSError:
            if time.time() > deadline:
                return
            time.sleep(0.005)
    time.sleep(hold_s)
    s.close()


def test_a_silent_stranger_does_not_stall_the_rendezvous():
    """Rank 0 drops a connection that does not announce a rank within two seconds and keeps accepting the real ranks."""
    nranks = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    boot_id = f"test-{uuid.uuid4().hex}-{os.getpid()}"
    stranger = ctx.Process(target=_silent_client, args=(boot_id, 12.0))
    procs = [ctx.Process(target=_worker, args=(boot_id, r, nranks, q)) for r in range(nranks)]
    stranger.start()
    for p in procs:
        p.start()
    results = [q.get(timeout=60) for _ in range(nranks)]
    for p in procs + [stranger]:
        p.join(timeout=30)
    for rank, st, msg in results:
        assert st == 0, f"rank {rank}: {msg}"
