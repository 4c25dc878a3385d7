"""GPU: the single-GPU kernel family against the fp32/fp64 CPU oracle (W3/W4 of SURVEY §8).

Matrix: dtype x op x n (incl. 0, 1, 2^k±1, non-vector-multiple), misaligned base pointers, NaN/inf,
every kernel variant x cluster size, determinism, 64-bit sizes. Tolerances follow the reference's self-check
(reduction.cpp:750-780: int exact, float 1e-8*n, double 1e-12) but relative to the magnitude of the data.
"""
import math

import pytest
import torch

import cuda_mpi_reductions_b200 as b2r

pytestmark = pytest.mark.gpu

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]
OPS = ["sum", "min", "max"]
SIZES = [0, 1, 2, 3, 31, 32, 33, 255, 256, 257, 1023, 4099, 65536, 65537, (1 << 20) - 1, (1 << 20) + 7, 1 << 24]
VARIANTS = ["ldg128", "ldg256", "tma"]


def check(got: torch.Tensor, x_cpu: torch.Tensor, op: str):
    want = b2r.reduce_oracle(x_cpu, op)
    g = got.cpu()
    n = x_cpu.numel()
    if x_cpu.dtype == torch.int32 or op != "sum":
        if want.is_floating_point() and math.isnan(want.item()):
            assert math.isnan(g.item())
        else:
            assert g.item() == want.item(), f"{x_cpu.dtype} {op} n={n}: got {g.item()} want {want.item()}"
        return
    mag = x_cpu.double().abs().sum().item() if n else 0.0
    if x_cpu.dtype == torch.float64:
        tol = 1e-15 * max(mag, 1e-300) * max(1.0, math.log2(max(n, 2)))
    elif x_cpu.dtype == torch.float32:
        tol = 6e-8 * max(mag, 1e-30) * max(1.0, math.log2(max(n, 2)))
    else:  # bf16: fp32 accumulate, ONE rounding to bf16 at the end
        exact = x_cpu.double().sum().item()
        tol = abs(exact) * 2.0**-8 + 6e-8 * mag * max(1.0, math.log2(max(n, 2)))
    assert abs(g.double().item() - want.double().item()) <= tol, (
        f"{x_cpu.dtype} {op} n={n}: got {g.item()} want {want.item()} tol {tol}")


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("op", OPS)
def test_matrix_default_dispatch(dtype, op):
    for n in SIZES:
        x = b2r.fill_host(n, dtype, seed=n, mode="signed")
        got = b2r.reduce(x.cuda(), op)
        check(got, x, op)


@pytest.mark.parametrize("variant", VARIANTS)
@pytest.mark.parametrize("cluster", [1, 2, 8])
@pytest.mark.parametrize("dtype", DTYPES)
def test_variants_and_clusters(variant, cluster, dtype):
    for n in [1, 4097, (1 << 18) + 3, (1 << 22) + 1]:
        x = b2r.fill_host(n, dtype, seed=11 * n, mode="signed")
        xd = x.cuda()
        for op in OPS:
            for threads in (128, 512):
                got = b2r.reduce(xd, op, cfg={"variant": variant, "cluster": cluster, "threads": threads, "ctas": 64})
                check(got, x, op)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("variant", VARIANTS)
def test_misaligned_base_pointers(dtype, variant):
    base = b2r.fill_host((1 << 16) + 64, dtype, seed=5, mode="signed")
    bd = base.cuda()
    for off in [1, 2, 3, 5, 7, 9, 15]:
        for n in [1, 17, 1000, (1 << 16)]:
            x = base[off:off + n]
            got = b2r.reduce(bd[off:off + n], "sum", cfg={"variant": variant})
            check(got, x, "sum")
            got = b2r.reduce(bd[off:off + n], "max", cfg={"variant": variant})
            check(got, x, "max")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
def test_nan_inf_negzero(dtype):
    n = (1 << 16) + 3
    x = b2r.fill_host(n, dtype, seed=9, mode="signed")
    for pos in [0, 1, n // 2, n - 1]:
        y = x.clone()
        y[pos] = float("nan")
        for op in OPS:
            assert math.isnan(b2r.reduce(y.cuda(), op).item()), (op, pos)
        z = x.clone()
        z[pos] = float("-inf")
        assert b2r.reduce(z.cuda(), "min").item() == -math.inf
        assert b2r.reduce(z.cuda(), "sum").item() == -math.inf
        z[pos] = float("inf")
        assert b2r.reduce(z.cuda(), "max").item() == math.inf
    w = torch.zeros(1000, dtype=dtype)
    w[3] = -0.0
    assert b2r.reduce(w.cuda(), "sum").item() == 0.0


def test_int_extremes_and_wrap():
    x = torch.full(((1 << 20) + 1,), 2**31 - 1, dtype=torch.int32)
    got = b2r.reduce(x.cuda(), "sum")
    check(got, x, "sum")
    x[12345] = -(2**31)
    assert b2r.reduce(x.cuda(), "min").item() == -(2**31)
    assert b2r.reduce(x.cuda(), "max").item() == 2**31 - 1
    full = b2r.fill_host(1 << 22, torch.int32, seed=3, mode="full")
    check(b2r.reduce(full.cuda(), "sum"), full, "sum")


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
@pytest.mark.parametrize("variant", VARIANTS)
def test_deterministic_bitwise(dtype, variant):
    x = b2r.fill_device((1 << 24) + 13, dtype, seed=1, mode="signed")
    first = b2r.reduce(x, "sum", cfg={"variant": variant}).clone()
    for _ in range(20):
        again = b2r.reduce(x, "sum", cfg={"variant": variant})
        assert torch.equal(first.reshape(1).view(torch.uint8).cpu(), again.reshape(1).view(torch.uint8).cpu())


def test_back_to_back_launches_reuse_ticket():
    # the ticket is re-armed by the last CTA; 200 stream-ordered launches with different sizes must all be right
    xs = [b2r.fill_device(n, torch.int32, seed=n, mode="full") for n in (1 << 20, (1 << 21) + 5, 77777, 1 << 22)]
    want = [b2r.reduce_oracle(x.cpu(), "sum").item() for x in xs]
    outs = []
    for i in range(200):
        outs.append(b2r.reduce(xs[i % 4], "sum"))
    torch.cuda.synchronize()
    for i, o in enumerate(outs):
        assert o.item() == want[i % 4]


def test_fill_device_matches_host():
    for dtype in DTYPES:
        for mode in ("sdk", "full", "signed"):
            d = b2r.fill_device(100003, dtype, seed=42, first_index=17, mode=mode)
            h = b2r.fill_host(100003, dtype, seed=42, first_index=17, mode=mode)
            assert torch.equal(d.cpu().view(torch.uint8), h.view(torch.uint8)), (dtype, mode)


@pytest.mark.parametrize("dtype", DTYPES)
def test_dot(dtype):
    n = (1 << 20) + 3
    a = b2r.fill_host(n, dtype, seed=1, mode="signed")
    c = b2r.fill_host(n, dtype, seed=2, mode="signed")
    got = b2r.dot(a.cuda(), c.cuda()).cpu()
    if dtype == torch.int32:
        want = (a.long() * c.long()).sum().item()
        want = (want + 2**31) % 2**32 - 2**31
        assert got.item() == want
    else:
        want = (a.double() * c.double()).sum().item()
        mag = (a.double() * c.double()).abs().sum().item()
        rel = {torch.float32: 1e-6, torch.float64: 1e-14, torch.bfloat16: 2.0**-7}[dtype]
        assert abs(got.double().item() - want) <= rel * max(mag if dtype != torch.bfloat16 else abs(want) + 1e-2 * mag, 1e-30)


def test_reduce_from_host_pipeline():
    for dtype in DTYPES:
        n = (3 << 21) + 11  # > 1 chunk for every dtype with the default 8 MiB chunk
        x = b2r.fill_host(n, dtype, seed=8, mode="signed", pin_memory=True)
        for op in OPS:
            got = b2r.reduce_from_host(x, op)
            check(got, x, op)


def test_trace_stamps():
    ws = b2r.Workspace()
    ws.trace(True)
    x = b2r.fill_device(1 << 24, torch.float32, seed=1)
    b2r.reduce(x, "sum", workspace=ws)
    torch.cuda.synchronize()
    t = b2r.trace_last(ws)
    assert t.shape[1] == 4 and t.shape[0] >= 148
    assert (t[:, 1] >= t[:, 0]).all() and (t[:, 2] >= t[:, 1]).all() and (t[:, 3] >= t[:, 2]).all()


@pytest.mark.slow
def test_two_pow_30_elements_and_beyond_int32_indexing():
    # 2^31 + 5 bf16 elements = 4 GiB: exceeds the reference's int indexing
    n = (1 << 31) + 5
    x = torch.ones(n, dtype=torch.bfloat16, device="cuda")
    x[-1] = 3.0
    assert b2r.reduce(x, "max").item() == 3.0
    assert b2r.reduce(x, "min").item() == 1.0
    s = b2r.reduce(x, "sum").item()
    assert abs(s - (n + 2)) <= (n + 2) * 2.0**-8
    del x
    y = b2r.fill_device(1 << 30, torch.int32, seed=2, mode="sdk")
    got = b2r.reduce(y, "sum").item()
    want = int(y.view(-1, 1 << 20).sum(dim=1, dtype=torch.int64).sum().item())
    assert got == (want + 2**31) % 2**32 - 2**31


def test_cuda_graph_capture_and_pdl():
    """No allocation / no sync in the launch path: K reductions (with programmatic dependent launch between them)
    can be captured in a CUDA graph and replayed."""
    xs = [b2r.fill_device(1 << 22, torch.int32, seed=i, mode="full") for i in range(4)]
    outs = [torch.zeros((), dtype=torch.int32, device="cuda") for _ in range(4)]
    want = [b2r.reduce_oracle(x.cpu(), "sum").item() for x in xs]
    ws = b2r.Workspace()
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for x, o in zip(xs, outs):
            b2r.reduce(x, "sum", out=o, workspace=ws, cfg=cfg)  # warm-up outside capture
    torch.cuda.synchronize()
    for o in outs:
        o.zero_()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=s):
        for x, o in zip(xs, outs):
            b2r.reduce(x, "sum", out=o, workspace=ws, cfg=cfg)
    for _ in range(3):
        g.replay()
    torch.cuda.synchronize()
    assert [o.item() for o in outs] == want


def test_pdl_back_to_back_is_correct():
    xs = [b2r.fill_device(n, torch.float64, seed=n, mode="signed") for n in (1 << 22, (1 << 23) + 5, 12345, 1 << 24)]
    want = [b2r.reduce(x, "sum").item() for x in xs]
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    ws = b2r.Workspace()
    outs = [torch.zeros((), dtype=torch.float64, device="cuda") for _ in range(400)]
    for i in range(400):
        b2r.reduce(xs[i % 4], "sum", out=outs[i], workspace=ws, cfg=cfg)
    torch.cuda.synchronize()
    # same values up to fp reassociation of a different launch geometry
    for i, o in enumerate(outs):
        assert abs(o.item() - want[i % 4]) <= 1e-9 * max(1.0, abs(want[i % 4]))
    # and the PDL path itself is deterministic
    assert len({outs[i].item() for i in range(0, 400, 4)}) == 1


def test_reduce_on_non_current_device_pointer_guard():
    # in == previous out: the overlap guard must drop PDL for that launch and still be right
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    ws = b2r.Workspace()
    x = b2r.fill_device(1 << 20, torch.float32, seed=3, mode="sdk")
    o1 = torch.zeros(1, dtype=torch.float32, device="cuda")
    b2r.reduce(x, "sum", out=o1, workspace=ws, cfg=cfg)
    o2 = b2r.reduce(o1, "sum", workspace=ws, cfg=cfg)
    assert o2.item() == o1.item()


@pytest.mark.parametrize("dtype", DTYPES)
def test_tma_single_cta_tile_follows_thread_count(dtype):
    """Regression: for inputs of exactly one planned tile (8 KiB with 160 planned threads) the single-CTA path lowers the
    thread count; the tile size has to follow or part of the tile is never read (found by bench_reduce --shmoo)."""
    esz = torch.empty((), dtype=dtype).element_size()
    for nbytes in (4096, 6144, 8192, 8192 + 16, 14336, 16384, 32768, 65536):
        n = nbytes // esz
        x = b2r.fill_host(n, dtype, seed=nbytes, mode="signed")
        for op in OPS:
            check(b2r.reduce(x.cuda(), op, cfg={"variant": "tma"}), x, op)
            check(b2r.reduce(x.cuda(), op, cfg={"variant": "tma", "pdl": 2}), x, op)
            check(b2r.reduce(x.cuda(), op, cfg={"variant": "tma", "pdl": 1}), x, op)


def test_safe_pdl_reads_an_input_written_by_the_preceding_kernel(monkeypatch):
    """with B2R_PDL=1 in the environment EVERY call site gets programmatic launches, including those whose
    input was written by the kernel right before them. The environment flavour therefore waits for that kernel before its
    first load. Producer -> reduce, many times, alternating two values so a stale read cannot pass by luck."""
    monkeypatch.setenv("B2R_PDL", "1")
    b2r.refresh_env()
    try:
        n = 1 << 24
        x = torch.empty(n, dtype=torch.int32, device="cuda")
        ws = b2r.Workspace()
        outs = [torch.zeros((), dtype=torch.int32, device="cuda") for _ in range(60)]
        assert b2r.plan(n, torch.int32)["pdl"] == 1
        for i in range(60):
            x.fill_(1 + (i % 2))  # the producer: a plain torch kernel, it never triggers programmatic completion
            b2r.reduce(x, "sum", out=outs[i], workspace=ws)
        torch.cuda.synchronize()
        for i, o in enumerate(outs):
            assert o.item() == n * (1 + (i % 2)), i
        # the same through dot -> reduce (dot's output scalar is the next input)
        a = torch.full((1 << 20,), 3.0, dtype=torch.float32, device="cuda")
        for i in range(20):
            d = b2r.dot(a, a, workspace=ws)
            r = b2r.reduce(d.reshape(1), "sum", workspace=ws)
            assert r.item() == 9.0 * (1 << 20)
    finally:
        monkeypatch.delenv("B2R_PDL")
        b2r.refresh_env()


@pytest.mark.parametrize("dtype", DTYPES)
def test_dynamic_tail_matches_oracle_and_is_reproducible(dtype):
    """The dynamically scheduled tail (tiles claimed with an atomic counter): right answers for every size class around the
    tile boundaries, misaligned bases included, and floating-point sums that do not depend on which CTA claimed which tile
    (per-tile partials folded in index order)."""
    esz = torch.empty((), dtype=dtype).element_size()
    ws = b2r.Workspace()
    for nbytes, cfg in (((8 << 20) + 48, {"dyn_pct": 25}), ((32 << 20) - 16, {"dyn_pct": 50, "threads": 256}), ((3 << 20) + 4 * esz, {"dyn_pct": 90, "ctas": 37, "threads": 128}),
                        ((16 << 20), {"dyn_pct": 25, "variant": "ldg256"}), ((16 << 20) + 1024, {"dyn_pct": 10, "unroll": 8})):
        n = nbytes // esz
        x = b2r.fill_host(n + 3, dtype, seed=nbytes, mode="signed")
        for off in (0, 3):
            xd = x.cuda()[off:off + n]
            for op in OPS:
                got = b2r.reduce(xd, op, workspace=ws, cfg=cfg)
                check(got, x[off:off + n], op)
            if dtype != torch.int32:
                first = b2r.reduce(xd, "sum", workspace=ws, cfg=cfg).cpu().reshape(1).view(torch.uint8)
                for _ in range(20):
                    again = b2r.reduce(xd, "sum", workspace=ws, cfg=cfg).cpu().reshape(1).view(torch.uint8)
                    assert torch.equal(first, again), (dtype, nbytes, cfg)


def test_dynamic_tail_is_what_the_default_plain_launch_runs():
    n = 1 << 24
    x = b2r.fill_device(n, torch.int32, seed=5, mode="sdk")
    assert b2r.plan(n, torch.int32, device=0)["dyn_pct"] > 0
    want = b2r.reduce_oracle(x.cpu(), "sum").item()
    ws = b2r.Workspace()
    for _ in range(30):  # the tile counter must be re-armed by every launch
        assert b2r.reduce(x, "sum", workspace=ws).item() == want
