"""CPU-only: the collective algorithms on the host loopback transport (threads as ranks, src/coll/host_algos.h) against
an element-wise oracle — MPI_Reduce semantics (mpi/reduce.c:76,90): recv[j] = op over ranks of send_r[j], valid at root."""
import pytest
import torch

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200 import _lib
from cuda_mpi_reductions_b200.parallel import host_reduce_threads

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]


def chain_oracle(xs, op, root):
    """Left fold from the farthest chain position down to the root (the ring algorithm's association order)."""
    P = len(xs)
    order = [xs[(root + (P - 1 - q)) % P] for q in range(P)]
    acc = order[0].clone()
    for x in order[1:]:  # pairwise: bf16 partials are rounded at every hop of the chain
        acc = elementwise_oracle([acc, x], op)
    return acc


def elementwise_oracle(xs, op):
    """Left-to-right combine in rank order; bf16 combined in fp32 and rounded once (the library's definition)."""
    if xs[0].dtype == torch.bfloat16:
        acc = xs[0].float()
        for x in xs[1:]:
            acc = acc + x.float() if op == "sum" else (torch.minimum(acc, x.float()) if op == "min" else torch.maximum(acc, x.float()))
        return acc.bfloat16()
    acc = xs[0].clone()
    for x in xs[1:]:
        b2r.combine_oracle(acc, x, op)
    return acc


@pytest.mark.parametrize("nranks", [1, 2, 3, 4, 7, 8])
@pytest.mark.parametrize("algo", ["oneshot", "twoshot", "tree", "push"])
@pytest.mark.parametrize("dtype", DTYPES)
def test_reduce_and_allreduce(nranks, algo, dtype):
    for count in (1, 5, 1023):
        xs = [b2r.fill_host(count, dtype, seed=100 * r + count, mode="signed") for r in range(nranks)]
        for op in ("sum", "min", "max"):
            want = elementwise_oracle(xs, op)
            for root in {0, nranks - 1, nranks // 2}:
                got = host_reduce_threads(xs, op, root, algo)
                if algo == "tree" and op == "sum" and dtype.is_floating_point:
                    tol = {torch.float32: 1e-5, torch.float64: 1e-13, torch.bfloat16: 0.07}[dtype]
                    assert torch.allclose(got.double(), want.double(), rtol=tol, atol=tol)  # tree order != rank order
                else:
                    assert torch.equal(got.view(torch.uint8), want.view(torch.uint8)), (nranks, algo, dtype, op, root, count)
            allr = host_reduce_threads(xs, op, -1, algo)
            for r in range(nranks):
                if algo == "tree" and op == "sum" and dtype.is_floating_point:
                    assert torch.allclose(allr[r].double(), want.double(), rtol=0.07, atol=0.07)
                else:
                    assert torch.equal(allr[r].view(torch.uint8), want.view(torch.uint8))
            # every rank holds the SAME bits after an allreduce
            for r in range(1, nranks):
                assert torch.equal(allr[r].view(torch.uint8), allr[0].view(torch.uint8))


def test_int_sum_wraps_across_ranks():
    xs = [torch.full((64,), 2**31 - 1, dtype=torch.int32) for _ in range(3)]
    got = host_reduce_threads(xs, "sum", 0, "twoshot")
    want = ((3 * (2**31 - 1) + 2**31) % 2**32) - 2**31
    assert (got == want).all()


POSTED_KNOBS = ("B2R_AUTO_POSTED_MAX_BYTES", "B2R_AUTO_POSTED_KEY_MAX_BYTES_8", "B2R_AUTO_POSTED_KEY_MAX_BYTES_4", "B2R_AUTO_POSTED_2GPU_MAX_BYTES")


def test_algo_selection_is_pure_and_sensible():
    sel = _lib.LIB.b2r_select_algo
    # small messages: the LL path (allreduce up to 64 KiB, reduce-to-root up to 256 KiB). Beyond: NVLS whenever the switch can do it (>= 4 GPUs, any dtype/op with a multimem form, int32
    # included since the 8-GPU measurement of round 2); otherwise one-shot for small messages and two-shot beyond
    assert sel(1, 1024, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(0, 8192, _lib.INT32, _lib.MIN, 8, 0) == _lib.ALGO_LL
    assert sel(1, 65536, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(1, 65540, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(0, 262144, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(0, 262148, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_POSTED  # mid-size reduce-to-root: stage and return, the root reduces via the switch
    assert sel(0, 1 << 20, _lib.INT32, _lib.MAX, 4, 1) == _lib.ALGO_POSTED
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.MIN, 8, 1) == _lib.ALGO_POSTED   # fp min/max travel as integer keys
    # native multimem forms: up to 8 MiB (then the barrier-framed NVLS kernel); keys: fp32 up to 32 MiB, fp64 up to 8 MiB (then two-shot)
    for dt, op in ((_lib.BFLOAT16, _lib.MAX), (_lib.FLOAT64, _lib.SUM), (_lib.INT32, _lib.SUM), (_lib.INT32, _lib.MIN), (_lib.FLOAT32, _lib.SUM)):
        assert sel(0, 8 << 20, dt, op, 8, 1) == _lib.ALGO_POSTED and sel(0, (8 << 20) + 8, dt, op, 8, 1) == _lib.ALGO_NVLS
    assert sel(0, 32 << 20, _lib.FLOAT32, _lib.MAX, 4, 1) == _lib.ALGO_POSTED and sel(0, (32 << 20) + 4, _lib.FLOAT32, _lib.MAX, 4, 1) == _lib.ALGO_TWOSHOT
    assert sel(0, 8 << 20, _lib.FLOAT64, _lib.MIN, 8, 1) == _lib.ALGO_POSTED and sel(0, 16 << 20, _lib.FLOAT64, _lib.MIN, 8, 1) == _lib.ALGO_TWOSHOT
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.MIN, 8, 0) == _lib.ALGO_TWOSHOT  # no switch, no posted
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.SUM, 2, 1) == _lib.ALGO_POSTED   # two GPUs: between the LL range and 2 MiB (the one-shot's barriers cost more than its pull saves)
    assert sel(0, 1 << 20, _lib.FLOAT64, _lib.MIN, 2, 1) == _lib.ALGO_POSTED
    assert sel(0, (2 << 20) + 4, _lib.FLOAT32, _lib.SUM, 2, 1) == _lib.ALGO_ONESHOT and sel(0, 1 << 20, _lib.FLOAT32, _lib.SUM, 2, 0) == _lib.ALGO_ONESHOT
    assert sel(1, 1 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS     # reduce-to-root only
    assert sel(0, 1 << 28, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(0, 1 << 30, _lib.FLOAT32, _lib.MIN, 8, 1) == _lib.ALGO_TWOSHOT
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.SUM, 2, 1) == _lib.ALGO_TWOSHOT  # NVLS loses at 2 GPUs
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 0) == _lib.ALGO_TWOSHOT
    assert sel(1, 65540, _lib.FLOAT32, _lib.SUM, 8, 0) == _lib.ALGO_TWOSHOT
    assert sel(1, 65540, _lib.FLOAT32, _lib.SUM, 2, 0) == _lib.ALGO_LL       # two GPUs: one peer per consumer, LL up to 256 KiB
    assert sel(1, 262148, _lib.FLOAT32, _lib.SUM, 2, 0) == _lib.ALGO_ONESHOT
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.MIN, 8, 1) == _lib.ALGO_TWOSHOT  # no multimem min for fp32
    assert sel(1, 1 << 30, _lib.FLOAT64, _lib.MAX, 8, 1) == _lib.ALGO_TWOSHOT
    assert sel(1, 1 << 30, _lib.BFLOAT16, _lib.MAX, 8, 1) == _lib.ALGO_NVLS
    assert sel(0, 1 << 28, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS  # 4-byte multimem ops measure like the 16-byte fp32 form ...
    assert sel(0, 1 << 30, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_RING  # ... except for reduce-to-root from 512 MiB up
    assert sel(1, 1 << 30, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(1, 1 << 28, _lib.INT32, _lib.MAX, 4, 1) == _lib.ALGO_NVLS
    assert sel(0, 64, _lib.FLOAT64, _lib.SUM, 2, 0) == _lib.ALGO_LL
    assert sel(0, 1 << 30, _lib.FLOAT64, _lib.SUM, 2, 1) == _lib.ALGO_ONESHOT  # 2-GPU reduce: the root pulls, at any size
    assert sel(0, 1 << 20, _lib.FLOAT64, _lib.SUM, 1, 0) == _lib.ALGO_ONESHOT


@pytest.mark.parametrize("nranks", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("dtype", DTYPES)
def test_ring_reduce_folds_in_chain_order(nranks, dtype):
    for count in (1, 7, 1025):
        xs = [b2r.fill_host(count, dtype, seed=50 * r + count, mode="signed") for r in range(nranks)]
        for op in ("sum", "min", "max"):
            for root in {0, nranks - 1, nranks // 2}:
                got = host_reduce_threads(xs, op, root, "ring")
                want = chain_oracle(xs, op, root)
                assert torch.equal(got.view(torch.uint8), want.view(torch.uint8)), (nranks, dtype, op, root, count)


def test_ll_path_limits_and_selection(monkeypatch):
    """Host-side rules of the LL path: size limit per communicator size, chosen by AUTO up to 8 KiB (B2R_AUTO_LL_MAX_BYTES),
    chosen under B2R_ALGO=7 for every message that fits, and its CPU restatement is the rank-order fold (== one-shot)."""
    lim = _lib.LIB.b2r_ll_max_bytes
    assert [lim(p) for p in (0, 1, 2, 4, 8, 16, 17)] == [0, 0, 32768, 32768, 26624, 13312, 0]
    for p in (2, 3, 4, 8, 16):  # two parities x P sources x (2 x payload) must fit between kOffLL and the push flag table at 1 MiB
        assert 2 * p * 2 * lim(p) <= (1 << 20) - 192 * 1024 and lim(p) % 16 == 0
    sel = _lib.LIB.b2r_select_algo
    monkeypatch.delenv("B2R_ALGO", raising=False)
    for all_ in (0, 1):
        assert sel(all_, 4, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_LL and sel(all_, 65536, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
        assert sel(all_, 262148, _lib.INT32, _lib.SUM, 8, 1) == (_lib.ALGO_NVLS if all_ else _lib.ALGO_POSTED)
    monkeypatch.setenv("B2R_AUTO_LL_MAX_BYTES", "0")          # allreduce crossover
    monkeypatch.setenv("B2R_AUTO_LL_REDUCE_MAX_BYTES", "0")   # reduce-to-root crossover
    assert sel(0, 4, _lib.INT32, _lib.SUM, 8, 0) == _lib.ALGO_ONESHOT and sel(1, 4, _lib.INT32, _lib.SUM, 8, 0) == _lib.ALGO_ONESHOT
    for var in POSTED_KNOBS:
        monkeypatch.setenv(var, "0")
    assert sel(0, 1 << 20, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS and sel(0, 1 << 20, _lib.FLOAT64, _lib.MAX, 8, 1) == _lib.ALGO_TWOSHOT
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS and sel(0, 1 << 20, _lib.FLOAT32, _lib.MIN, 8, 1) == _lib.ALGO_TWOSHOT
    for var in POSTED_KNOBS:
        monkeypatch.delenv(var)
    monkeypatch.delenv("B2R_AUTO_LL_MAX_BYTES")
    monkeypatch.delenv("B2R_AUTO_LL_REDUCE_MAX_BYTES")
    monkeypatch.setenv("B2R_ALGO", str(_lib.ALGO_LL))
    assert sel(0, 262144, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(1, 262144, _lib.INT32, _lib.SUM, 8, 1) == _lib.ALGO_LL      # forced: every message that fits, allreduce too
    assert sel(1, 4, _lib.FLOAT64, _lib.MAX, 2, 0) == _lib.ALGO_LL
    assert sel(0, 262148, _lib.INT32, _lib.SUM, 8, 0) == _lib.ALGO_TWOSHOT  # too big: the normal choice
    assert sel(1, 1 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    monkeypatch.delenv("B2R_ALGO")
    xs = [b2r.fill_host(1001, torch.bfloat16, seed=9 * r, mode="signed") for r in range(5)]
    for root in (-1, 0, 3):
        a = host_reduce_threads(xs, "sum", root, "ll")
        b = host_reduce_threads(xs, "sum", root, "oneshot")
        assert torch.equal(a.view(torch.uint8), b.view(torch.uint8))
    name = _lib.LIB.b2r_algo_name
    name.restype = __import__("ctypes").c_char_p
    assert name(_lib.ALGO_LL) == b"ll" and name(_lib.ALGO_RING) == b"ring"


@pytest.mark.parametrize("nranks", [1, 2, 3, 5, 8])
@pytest.mark.parametrize("dtype", DTYPES)
def test_ring_allreduce_is_one_chain_per_slice(nranks, dtype):
    """Ring allreduce = P chains at once: chain c ends at rank c and folds slice c (cut at 16-byte vectors, the
    sub-vector tail rides with the last chain); every rank ends up with the same bits."""
    per_vec = 16 // torch.empty(0, dtype=dtype).element_size()
    for count in (1, per_vec - 1, per_vec, 7 * per_vec + 1, 1025):
        xs = [b2r.fill_host(count, dtype, seed=13 * r + count, mode="signed") for r in range(nranks)]
        nvec = count // per_vec
        for op in ("sum", "min", "max"):
            rows = host_reduce_threads(xs, op, -1, "ring")
            want = torch.empty_like(xs[0])
            for c in range(nranks):
                lo = nvec * c // nranks * per_vec
                hi = count if c == nranks - 1 else nvec * (c + 1) // nranks * per_vec
                if hi > lo:
                    want[lo:hi] = chain_oracle([x[lo:hi] for x in xs], op, c)
            for r in range(nranks):
                assert torch.equal(rows[r].view(torch.uint8), want.view(torch.uint8)), (nranks, dtype, op, count, r)


def test_auto_crossover_switches(monkeypatch):
    """B2R_AUTO_LL_MAX_BYTES (default 8192), B2R_AUTO_RING_MIN_MB and the three B2R_AUTO_PUSH_*_MB crossovers move AUTO's
    crossovers without a rebuild; unsetting them restores the measured defaults."""
    sel = _lib.LIB.b2r_select_algo
    for var in POSTED_KNOBS + ("B2R_ALGO", "B2R_AUTO_LL_MAX_BYTES", "B2R_AUTO_LL_REDUCE_MAX_BYTES", "B2R_AUTO_RING_MIN_MB", "B2R_AUTO_PUSH_REDUCE_MB", "B2R_AUTO_PUSH_ALLREDUCE_MB", "B2R_AUTO_PUSH_REDUCE_NVLS_MB"):
        monkeypatch.delenv(var, raising=False)
    before = [sel(a, n, d, _lib.SUM, p, 1) for a in (0, 1) for n in (64, 8192, 1 << 20, 1 << 28) for d in (_lib.INT32, _lib.FLOAT32) for p in (2, 8)]
    assert sel(0, 8192, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(1, 64, _lib.INT32, _lib.MIN, 2, 0) == _lib.ALGO_LL
    monkeypatch.setenv("B2R_AUTO_LL_REDUCE_MAX_BYTES", "8192")
    assert sel(0, 8193, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_POSTED  # what takes over from LL for reduce-to-root
    assert sel(1, 8193, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL     # the allreduce crossover is a separate knob
    monkeypatch.setenv("B2R_AUTO_LL_REDUCE_MAX_BYTES", str(1 << 30))  # capped by what the slots hold (256 KiB)
    assert sel(0, 262144, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_LL
    assert sel(0, 262148, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_POSTED
    monkeypatch.setenv("B2R_AUTO_POSTED_MAX_BYTES", str(1 << 20))
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_POSTED and sel(0, (1 << 20) + 4, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    monkeypatch.delenv("B2R_AUTO_POSTED_MAX_BYTES")
    monkeypatch.delenv("B2R_AUTO_LL_REDUCE_MAX_BYTES")
    monkeypatch.setenv("B2R_AUTO_RING_MIN_MB", "64")
    assert sel(0, 64 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_RING
    assert sel(0, 64 << 20, _lib.INT32, _lib.MAX, 4, 0) == _lib.ALGO_RING
    assert sel(0, (64 << 20) - 4, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS  # allreduce: the ring is reduce-only
    assert sel(0, 1 << 30, _lib.FLOAT64, _lib.SUM, 2, 1) == _lib.ALGO_ONESHOT  # two GPUs: the root pulls
    monkeypatch.delenv("B2R_AUTO_RING_MIN_MB")
    monkeypatch.setenv("B2R_AUTO_PUSH_REDUCE_MB", "32")
    monkeypatch.setenv("B2R_AUTO_PUSH_ALLREDUCE_MB", "128")
    monkeypatch.setenv("B2R_AUTO_PUSH_REDUCE_NVLS_MB", "0")
    assert sel(0, 32 << 20, _lib.FLOAT64, _lib.MIN, 8, 1) == _lib.ALGO_PUSH      # no multimem form: stores instead of pulls
    assert sel(0, (32 << 20) - 8, _lib.FLOAT64, _lib.MIN, 8, 1) == _lib.ALGO_TWOSHOT
    assert sel(0, (32 << 20) - 8, _lib.FLOAT32, _lib.MIN, 8, 1) == _lib.ALGO_POSTED   # below the push crossover: the usual choice
    assert sel(1, 32 << 20, _lib.FLOAT32, _lib.MAX, 8, 1) == _lib.ALGO_TWOSHOT   # allreduce has its own crossover
    assert sel(1, 128 << 20, _lib.FLOAT32, _lib.MAX, 8, 1) == _lib.ALGO_PUSH
    assert sel(0, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS       # the switch keeps what it can do ...
    monkeypatch.setenv("B2R_AUTO_PUSH_REDUCE_NVLS_MB", "64")
    assert sel(0, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_PUSH       # ... unless told otherwise: reduce-to-root only
    assert sel(0, 1 << 28, _lib.INT32, _lib.MAX, 8, 1) == _lib.ALGO_PUSH
    assert sel(0, 1 << 20, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_POSTED
    assert sel(1, 1 << 30, _lib.FLOAT32, _lib.SUM, 8, 1) == _lib.ALGO_NVLS
    assert sel(0, 1 << 30, _lib.FLOAT32, _lib.SUM, 2, 1) == _lib.ALGO_ONESHOT    # two GPUs: one worker would do all the work
    for var in ("B2R_AUTO_PUSH_REDUCE_MB", "B2R_AUTO_PUSH_ALLREDUCE_MB", "B2R_AUTO_PUSH_REDUCE_NVLS_MB"):
        monkeypatch.delenv(var)
    after = [sel(a, n, d, _lib.SUM, p, 1) for a in (0, 1) for n in (64, 8192, 1 << 20, 1 << 28) for d in (_lib.INT32, _lib.FLOAT32) for p in (2, 8)]
    assert before == after


def test_ll_big_region_arithmetic():
    """Sizes of the large LL instance (coll_ctx.h), restated: 64 CTAs x 256 vectors, acknowledgement words at the
    end of the region, two parities x P sources x slot, payload = half a slot."""
    ack = 64 * 16 * 4
    for region_mb, P, want in ((16, 8, 256 << 10), (16, 16, 262016), (16, 2, 256 << 10), (4, 8, 130944), (1, 16, 16256)):
        region = region_mb << 20
        slot = ((region - ack) // (2 * P)) & ~255
        assert min(slot // 2, 64 * 256 * 16) == want, (region_mb, P)
        assert 2 * P * slot + ack <= region
