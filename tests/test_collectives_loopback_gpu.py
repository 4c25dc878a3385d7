"""GPU (one device is enough): the cross-rank collectives on the LOOPBACK transport — P virtual ranks on cuda:0, all
ranks of a collective fused into one cooperative launch — against the host algorithms that combine in the same order.
SURVEY §4 "tests/dist": every root, in place and out of place, counts not divisible by P x vector width, symmetric and
staged buffers, repeated back-to-back calls (flag reuse), reduce + allreduce + scalar-sharded, timeout / fault injection."""
import os

import pytest
import torch

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200 import _lib
from cuda_mpi_reductions_b200.parallel import Communicator, host_reduce_threads

pytestmark = pytest.mark.gpu

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]
ALGOS = ["oneshot", "twoshot", "tree", "push"]
COUNTS = [1, 3, 7, 1000, 4099, 65537, (1 << 20) + 5]


@pytest.fixture(scope="module", params=[2, 3, 4, 8])
def comms(request):
    cs = Communicator.loopback(request.param, device=0, heap_bytes=64 << 20)
    yield cs
    for c in cs:
        c.close()


def bits(t):
    return t.reshape(-1).view(torch.uint8)


@pytest.mark.parametrize("algo", ALGOS)
@pytest.mark.parametrize("dtype", DTYPES)
def test_allreduce(comms, algo, dtype):
    P = len(comms)
    for count in COUNTS:
        xs_h = [b2r.fill_host(count, dtype, seed=31 * r + count, mode="signed") for r in range(P)]
        xs = [x.cuda() for x in xs_h]
        for op in ("sum", "min", "max"):
            want = host_reduce_threads(xs_h, op, -1, algo)[0]
            outs = Communicator.group_allreduce(comms, xs, op, algo=algo)
            Communicator.group_check(comms)
            for r in range(P):
                assert torch.equal(bits(outs[r].cpu()), bits(want)), (P, algo, dtype, op, count, r)


@pytest.mark.parametrize("algo", ALGOS + ["ring"])
@pytest.mark.parametrize("dtype", [torch.int32, torch.float64, torch.bfloat16])
def test_reduce_every_root(comms, algo, dtype):
    P = len(comms)
    for count in (1, 1000, 65537, (1 << 21) + 9):
        xs_h = [b2r.fill_host(count, dtype, seed=17 * r + count, mode="signed") for r in range(P)]
        xs = [x.cuda() for x in xs_h]
        for root in range(P):
            want = host_reduce_threads(xs_h, "sum", root, algo)
            outs = Communicator.group_reduce(comms, xs, "sum", root=root, algo=algo)
            Communicator.group_check(comms)
            assert torch.equal(bits(outs[root].cpu()), bits(want)), (P, algo, dtype, count, root)
            for r in range(P):  # send buffers are never modified
                assert torch.equal(bits(xs[r].cpu()), bits(xs_h[r]))


@pytest.mark.parametrize("algo", ALGOS)
def test_symmetric_buffers_in_place_and_offsets(comms, algo):
    """Zero-copy path: buffers from the symmetric heap, deliberately at DIFFERENT offsets on different ranks (the
    kernels exchange offsets in their start barrier), in place and out of place."""
    P = len(comms)
    count = 50001
    for c in comms:
        c.symm_free_all()
    sends, recvs = [], []
    for r, c in enumerate(comms):
        c.symm_empty(64 * (r + 1), torch.float32)  # skew the offsets
        sends.append(c.symm_empty(count, torch.float32))
        recvs.append(c.symm_empty(count, torch.float32))
    xs_h = [b2r.fill_host(count, torch.float32, seed=5 + r, mode="signed") for r in range(P)]
    want = host_reduce_threads(xs_h, "sum", -1, algo)[0]
    for s, x in zip(sends, xs_h):
        s.copy_(x)
    outs = Communicator.group_allreduce(comms, sends, "sum", algo=algo, outs=recvs)
    Communicator.group_check(comms)
    for r in range(P):
        assert outs[r].data_ptr() == recvs[r].data_ptr()
        assert torch.equal(bits(recvs[r].cpu()), bits(want))
    # in place
    outs = Communicator.group_allreduce(comms, sends, "sum", algo=algo, outs=sends)
    Communicator.group_check(comms)
    for r in range(P):
        assert torch.equal(bits(sends[r].cpu()), bits(want))


def test_back_to_back_small_collectives_reuse_flags(comms):
    P = len(comms)
    xs_h = [b2r.fill_host(257, torch.int32, seed=r, mode="full") for r in range(P)]
    xs = [x.cuda() for x in xs_h]
    want = host_reduce_threads(xs_h, "sum", -1, "oneshot")[0]
    outs = None
    for i in range(300):
        algo = ALGOS[i % len(ALGOS)]
        outs = Communicator.group_allreduce(comms, xs, "sum", algo=algo)
    Communicator.group_check(comms)
    for o in outs:
        assert torch.equal(o.cpu(), want)


def test_large_message_is_chunked_through_staging():
    cs = Communicator.loopback(2, device=0, heap_bytes=1 << 20)
    try:
        count = (70 << 20) // 4 + 3  # > one 64 MiB staging region -> 2 launches
        xs = [b2r.fill_device(count, torch.float32, seed=r, mode="signed") for r in range(2)]
        outs = Communicator.group_allreduce(cs, xs, "sum", algo="twoshot")
        Communicator.group_check(cs)
        want = xs[0] + xs[1]
        assert torch.equal(outs[0], want) and torch.equal(outs[1], want)
        red = Communicator.group_reduce(cs, xs, "max", root=1, algo="oneshot")
        Communicator.group_check(cs)
        assert torch.equal(red[1], torch.maximum(xs[0], xs[1]))
    finally:
        for c in cs:
            c.close()


@pytest.mark.parametrize("dtype", DTYPES)
def test_scalar_sharded_reduce(comms, dtype):
    P = len(comms)
    for count in (1, 1000, (1 << 20) + 3):
        xs_h = [b2r.fill_host(count, dtype, seed=3 * r + count, mode="sdk") for r in range(P)]
        xs = [x.cuda() for x in xs_h]
        whole = torch.cat(xs_h)
        for op in ("sum", "min", "max"):
            want = b2r.reduce_oracle(whole, op)
            for root, allflag in ((0, False), (P - 1, False), (P // 2, True)):
                outs = Communicator.group_reduce_scalar(comms, xs, op, root=root, all=allflag)
                Communicator.group_check(comms)
                holders = range(P) if allflag else [root]
                for r in holders:
                    got = outs[r].cpu()
                    if dtype == torch.int32 or op != "sum":
                        assert got.item() == want.item(), (P, dtype, op, count, root, r)
                    else:
                        tol = {torch.float32: 1e-5, torch.float64: 1e-13, torch.bfloat16: 2.0**-7}[dtype]
                        assert abs(got.double().item() - want.double().item()) <= tol * max(1e-30, abs(want.double().item()))
    # many back-to-back calls: mailbox slots are double-buffered by call parity with root acks
    xs = [b2r.fill_device(4096, torch.int32, seed=r, mode="full") for r in range(P)]
    want = b2r.reduce_oracle(torch.cat([x.cpu() for x in xs]), "sum").item()
    for i in range(100):
        outs = Communicator.group_reduce_scalar(comms, xs, "sum", root=i % P, all=False)
        assert outs[i % P] is not None
    Communicator.group_check(comms)
    assert outs[99 % P].item() == want


def test_barrier_and_pending_semantics(comms):
    P = len(comms)
    for c in comms:
        c.barrier()
    Communicator.group_check(comms)
    # a rank may not issue two collectives before the group completed the first
    x = torch.zeros(8, device="cuda")
    if P > 1:
        comms[0].allreduce(x)
        with pytest.raises(b2r.B2RError):
            comms[0].allreduce(x)
        for c in comms[1:]:
            c.allreduce(x)
        Communicator.group_check(comms)


def test_mismatched_arguments_are_rejected(comms):
    P = len(comms)
    if P < 2:
        pytest.skip("needs 2 ranks")
    x = torch.zeros(8, device="cuda")
    y = torch.zeros(9, device="cuda")
    for c in comms[:-1]:
        c.allreduce(x)
    with pytest.raises(b2r.B2RError):
        comms[-1].allreduce(y)


def test_timeout_instead_of_hang(monkeypatch):
    """Fault injection: rank 1 never signals; its peers must give up after the deadline, flag the error
    word, and the kernel must terminate (no hang)."""
    monkeypatch.setenv("B2R_FAULT_INJECT", "rank:1,skip_signal:1")
    cs = Communicator.loopback(3, device=0, heap_bytes=1 << 20, timeout_ms=200)
    monkeypatch.delenv("B2R_FAULT_INJECT")
    try:
        xs = [torch.ones(1000, device="cuda") for _ in range(3)]
        Communicator.group_allreduce(cs, xs, "sum", algo="oneshot")
        statuses = []
        for c in cs:
            try:
                c.check()
                statuses.append(_lib.OK)
            except b2r.B2RError as e:
                statuses.append(e.status)
        assert statuses[0] == _lib.ERR_TIMEOUT and statuses[2] == _lib.ERR_TIMEOUT
    finally:
        for c in cs:
            c.close()


def test_delay_injection_is_tolerated(monkeypatch):
    monkeypatch.setenv("B2R_FAULT_INJECT", "rank:0,delay_us:50000")
    cs = Communicator.loopback(2, device=0, heap_bytes=1 << 20, timeout_ms=2000)
    monkeypatch.delenv("B2R_FAULT_INJECT")
    try:
        xs = [torch.full((100,), float(r + 1), device="cuda") for r in range(2)]
        outs = Communicator.group_allreduce(cs, xs, "sum", algo="twoshot")
        Communicator.group_check(cs)
        assert (outs[0] == 3).all() and (outs[1] == 3).all()
    finally:
        for c in cs:
            c.close()


def test_weak_payload_flavour_in_a_fresh_process():
    """B2R_WEAK_DATA=1 switches the payload loads/stores of the collective kernels to weak accesses (a per-process,
    per-device constant), so it is exercised in a subprocess: every algorithm must still match the host algorithms."""
    import subprocess
    import sys

    code = r'''
import torch, cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200.parallel import Communicator, host_reduce_threads
cs = Communicator.loopback(4, device=0, heap_bytes=32 << 20)
for dtype in (torch.int32, torch.float64, torch.bfloat16):
    for count in (5, 70001, (1 << 20) + 3):
        xh = [b2r.fill_host(count, dtype, seed=9 * r + count, mode="signed") for r in range(4)]
        xs = [x.cuda() for x in xh]
        for algo in ("oneshot", "twoshot", "tree", "push"):
            want = host_reduce_threads(xh, "sum", -1, algo)[0]
            outs = Communicator.group_allreduce(cs, xs, "sum", algo=algo)
            Communicator.group_check(cs)
            assert all(torch.equal(o.cpu().view(torch.uint8), want.view(torch.uint8)) for o in outs), (dtype, count, algo)
            red = Communicator.group_reduce(cs, xs, "max", root=2, algo=algo)
            Communicator.group_check(cs)
            assert torch.equal(red[2].cpu().view(torch.uint8), host_reduce_threads(xh, "max", 2, algo).view(torch.uint8))
print("WEAK_OK")
'''
    env = dict(os.environ, B2R_WEAK_DATA="1")
    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    res = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300, env=env, cwd=repo)
    assert res.returncode == 0 and "WEAK_OK" in res.stdout, res.stdout[-1000:] + res.stderr[-2000:]


def test_single_rank_communicator_degenerates_to_a_copy():
    (c,) = Communicator.loopback(1, device=0, heap_bytes=1 << 20)
    try:
        x = b2r.fill_device(1000, torch.float32, seed=1, mode="signed")
        assert torch.equal(c.allreduce(x, "sum"), x)
        assert torch.equal(c.reduce(x, "max", root=0), x)
        got = c.reduce_scalar_sharded(x, "sum")
        c.check()
        assert abs(got.item() - x.double().sum().item()) < 1e-3
        c.barrier()
        assert c.select_algo(1 << 20, torch.float32) == "oneshot"
    finally:
        c.close()


def test_ring_is_reduce_only(comms):
    x = torch.zeros(64, device="cuda")
    with pytest.raises(b2r.B2RError):
        for c in comms:
            c.allreduce(x, "sum", algo="ring")
    # the failed group call must not leave a half-recorded collective behind
    outs = Communicator.group_allreduce(comms, [x] * len(comms), "sum", algo="twoshot")
    Communicator.group_check(comms)
    assert all(float(o.sum()) == 0.0 for o in outs)
