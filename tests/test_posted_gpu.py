"""GPU (loopback transport): the posted reduce-to-root (`algo="posted"`, src/coll/coll_posted.cuh).

On a loopback group there is no switch, so the root folds the staged copies with peer loads — the protocol under test is the
same one the multi-GPU path runs (fixed chunk ownership, per-CTA ready words, slot parity, acknowledgement of the root of the
call two back); the multimem instruction itself is exercised by tests/dist_worker.py on real GPUs. Results must be
bit-identical to the rank-order fold of the host restatement."""
import pytest
import torch

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200.parallel import Communicator, host_reduce_threads

pytestmark = [pytest.mark.gpu]

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]
CHUNK_BYTES = 512 * 16  # kPostedChunkVecs vectors
GRID = 128              # kPostedCtas on real GPUs; a loopback group of P ranks uses min(128, SMs / P) CTAs per rank, so the
                        # sizes below cross "every CTA has one chunk" at different points for different P


@pytest.fixture(scope="module", params=[2, 3, 8])
def comms(request):
    cs = Communicator.loopback(request.param, device=0, heap_bytes=16 << 20)
    yield cs
    for c in cs:
        c.close()


def bits(t):
    return t.reshape(-1).view(torch.uint8)


def counts_for(dtype):
    esz = torch.empty(0, dtype=dtype).element_size()
    per_vec = 16 // esz
    chunk = CHUNK_BYTES // esz
    return sorted({1, per_vec - 1, per_vec, per_vec + 1, chunk - 1, chunk, chunk + per_vec + 1, 3 * chunk + 5, GRID * chunk, GRID * chunk + 1,
                   (GRID + 1) * chunk + per_vec - 1, 2 * GRID * chunk + 3 * chunk + 7} - {0})


@pytest.mark.parametrize("dtype", DTYPES)
def test_posted_reduce_every_size_and_op(comms, dtype):
    P = len(comms)
    for i, count in enumerate(counts_for(dtype)):
        xs_h = [b2r.fill_host(count, dtype, seed=17 * r + count, mode="signed") for r in range(P)]
        xs = [x.cuda() for x in xs_h]
        for j, op in enumerate(("sum", "min", "max")):
            root = (i + j) % P
            want = host_reduce_threads(xs_h, op, root, "posted")
            outs = Communicator.group_reduce(comms, xs, op, root=root, algo="posted")
            Communicator.group_check(comms)
            assert torch.equal(bits(outs[root].cpu()), bits(want)), (P, dtype, op, count, root)
        for r in range(P):  # send buffers are never modified
            assert torch.equal(bits(xs[r].cpu()), bits(xs_h[r]))


def test_posted_in_place_at_the_root_and_equal_to_oneshot(comms):
    P = len(comms)
    count = 5 * (CHUNK_BYTES // 4) + 3
    xs_h = [b2r.fill_host(count, torch.float32, seed=5 + r, mode="signed") for r in range(P)]
    for root in (0, P - 1):
        ref = Communicator.group_reduce(comms, [x.cuda() for x in xs_h], "sum", root=root, algo="oneshot")
        xs = [x.cuda() for x in xs_h]
        outs = [xs[r] if r == root else None for r in range(P)]
        got = Communicator.group_reduce(comms, xs, "sum", root=root, algo="posted", outs=outs)
        Communicator.group_check(comms)
        assert got[root].data_ptr() == xs[root].data_ptr()
        assert torch.equal(bits(xs[root].cpu()), bits(ref[root].cpu()))
        for r in range(P):
            if r != root:
                assert torch.equal(bits(xs[r].cpu()), bits(xs_h[r]))


def test_posted_back_to_back_mixed_sequence(comms):
    """400 calls with no host synchronisation in between: roots rotate with a period that is coprime to the slot parity, sizes
    change (stale data of longer messages stays in the slots), and other algorithms — including push and the tree, which use
    the same work region, and LL — are interleaved."""
    P = len(comms)
    dtype = torch.int32
    sizes = [3, 4099, CHUNK_BYTES // 4 * 2 + 1, CHUNK_BYTES // 4 * (GRID + 3)]
    data = {n: [b2r.fill_host(n, dtype, seed=13 * r + n, mode="full") for r in range(P)] for n in sizes}
    dev = {n: [x.cuda() for x in xs] for n, xs in data.items()}
    want = {n: host_reduce_threads(xs, "sum", -1, "oneshot")[0] for n, xs in data.items()}
    others = ["twoshot", "push", "tree", "ll"]
    checks = []
    for i in range(400):
        n = sizes[(i * 3) % len(sizes)]
        root = (i * 5 + i // 7) % P
        if i % 9 == 8:
            algo = others[(i // 9) % len(others)]
            if algo == "ll" and n * 4 > comms[0].ll_max_bytes:
                algo = "twoshot"
            outs = Communicator.group_reduce(comms, dev[n], "sum", root=root, algo=algo)
        else:
            outs = Communicator.group_reduce(comms, dev[n], "sum", root=root, algo="posted")
        checks.append((n, root, outs))
        if len(checks) > 30:
            torch.cuda.synchronize()
            for n2, root2, outs2 in checks:
                assert torch.equal(outs2[root2].cpu(), want[n2]), (i, n2, root2)
            checks = []
    Communicator.group_check(comms)
    for n2, root2, outs2 in checks:
        assert torch.equal(outs2[root2].cpu(), want[n2])


def test_posted_float_min_max_special_values(comms):
    """The pull path folds with a < b ? a : b like every other kernel; the values that the integer-key form of the switch path
    has to order correctly (infinities, denormals, both zeros apart) go through here as well so that both paths see them."""
    P = len(comms)
    for dtype in (torch.float32, torch.float64):
        tiny = torch.finfo(dtype).tiny
        special = torch.tensor([float("inf"), -float("inf"), tiny / 4, -tiny / 4, 1.0, -1.0, torch.finfo(dtype).max, torch.finfo(dtype).min], dtype=dtype)
        xs_h = [special.roll(r).repeat(33) for r in range(P)]
        xs = [x.cuda() for x in xs_h]
        for op in ("min", "max"):
            want = host_reduce_threads(xs_h, op, 0, "posted")
            outs = Communicator.group_reduce(comms, xs, op, root=0, algo="posted")
            Communicator.group_check(comms)
            assert torch.equal(bits(outs[0].cpu()), bits(want)), (dtype, op)


def test_posted_limits_and_misaligned_views(comms):
    P = len(comms)
    assert comms[0].posted_max_bytes == 64 << 20
    xs = [torch.zeros(16, dtype=torch.int32, device="cuda") for _ in range(P)]
    with pytest.raises(b2r.B2RError, match="reduce-to-root only"):
        Communicator.group_allreduce(comms, xs, "sum", algo="posted")
    base_h = [b2r.fill_host(6000, torch.int32, seed=70 + r, mode="full") for r in range(P)]
    base = [x.cuda() for x in base_h]
    want = host_reduce_threads([x[1:5001].contiguous() for x in base_h], "sum", 1 % P, "posted")
    store = torch.zeros(5010, dtype=torch.int32, device="cuda")
    outs = [store[3:5003] if r == 1 % P else None for r in range(P)]
    got = Communicator.group_reduce(comms, [b[1:5001] for b in base], "sum", root=1 % P, algo="posted", outs=outs)
    Communicator.group_check(comms)
    assert torch.equal(got[1 % P].cpu(), want)
    assert int(store[:3].abs().sum()) == 0 and int(store[5003:].abs().sum()) == 0
