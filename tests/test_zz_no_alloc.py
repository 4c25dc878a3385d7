"""GPU: b2r_reduce_local_noalloc — the reduction for callers that cannot allocate device memory (a process whose GPU
memory belongs to another allocator, a loop that may not run device-synchronising calls such as cudaMalloc/cudaFree, a
hook whose contract forbids allocation). Its scratch is a static __device__ array inside libb2r.so instead of a
b2r_workspace.

The entry point is called through the C ABI (ctypes on libb2r.so, no Python wrapper in between) and compared with the
existing entry point b2r_reduce_local on two calls in a row: the second call finds the array as the first one left it,
which is exactly the situation the ticket clearing exists for. This file sorts last on purpose, so a run that stops at
the first failure has already exercised everything else; it must also pass alone in a fresh process (first call = the
one-time set-up on the device).
"""
import ctypes as C
import math
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), "cuda_mpi_reductions_b200", "lib", "libb2r.so")

INT32, FLOAT32, FLOAT64, BFLOAT16 = range(4)
SUM, MIN, MAX = range(3)
DTYPES = {torch.int32: INT32, torch.float32: FLOAT32, torch.float64: FLOAT64, torch.bfloat16: BFLOAT16}
MAX_PARTIALS = 1024  # B2R_NOALLOC_MAX_PARTIALS in include/b2r.h


class LaunchCfg(C.Structure):
    _fields_ = [(n, C.c_int) for n in ("variant", "threads", "ctas", "cluster", "stages", "tile_bytes", "unroll", "pdl", "dyn_pct")]


@pytest.fixture(scope="module")
def lib():
    lib = C.CDLL(LIB_PATH)
    u64, i, vp = C.c_uint64, C.c_int, C.c_void_p
    lib.b2r_reduce_local_noalloc.argtypes = [vp, vp, u64, i, i, vp, C.POINTER(LaunchCfg)]
    lib.b2r_reduce_local_noalloc.restype = i
    lib.b2r_reduce_local.argtypes = [vp, vp, u64, i, i, vp, vp]
    lib.b2r_reduce_local.restype = i
    lib.b2r_reduce_local_ex.argtypes = [vp, vp, u64, i, i, vp, vp, C.POINTER(LaunchCfg)]
    lib.b2r_reduce_local_ex.restype = i
    lib.b2r_refresh_env.restype = None
    lib.b2r_workspace_create.argtypes = [C.POINTER(vp), i]
    lib.b2r_workspace_create.restype = i
    lib.b2r_workspace_destroy.argtypes = [vp]
    lib.b2r_workspace_destroy.restype = i
    lib.b2r_last_error_string.restype = C.c_char_p
    return lib


def make_input(n, dtype, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if dtype == torch.int32:
        return torch.randint(-2**31, 2**31 - 1, (n,), dtype=torch.int32, generator=g).cuda()
    return (torch.rand(n, generator=g, dtype=torch.float64) * 2 - 1).to(dtype).cuda()


def agree(a: torch.Tensor, b: torch.Tensor, dtype, op, n, mag):
    """Same tolerance as tests/test_reduce_local_gpu.py: integers and min/max exact; floating-point sums within the
    accumulation bound, which scales with the sum of magnitudes `mag` (the two entry points may use different launch
    geometries — static split here, dynamically scheduled tail there — so their partial sums associate differently)."""
    x, y = a.cpu().double().item(), b.cpu().double().item()
    if dtype == torch.int32 or op != SUM:
        return x == y or (math.isnan(x) and math.isnan(y))
    if dtype == torch.bfloat16:  # fp32 accumulation, one rounding to bf16 at the end
        return abs(x - y) <= 2.0**-7 * max(abs(x), abs(y)) + 6e-8 * mag * math.log2(max(n, 2))
    rel = {torch.float32: 6e-8, torch.float64: 1e-15}[dtype]
    return abs(x - y) <= 2 * rel * max(mag, 1e-30) * max(1.0, math.log2(max(n, 2)))


@pytest.mark.parametrize("dtype", list(DTYPES))
def test_noalloc_matches_workspace_entry_point_twice_in_a_row(lib, dtype):
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    ws = C.c_void_p()
    assert lib.b2r_workspace_create(C.byref(ws), 0) == 0
    try:
        for op in (SUM, MIN, MAX):
            # two different inputs back to back: the second call starts from the scratch contents of the first
            for n, seed in (((1 << 22) + 13, 1), (777_777, 2)):
                x = make_input(n, dtype, seed)
                want = torch.zeros((), dtype=dtype, device="cuda")
                got = torch.zeros((), dtype=dtype, device="cuda")
                assert lib.b2r_reduce_local(x.data_ptr(), want.data_ptr(), n, DTYPES[dtype], op, ws, stream) == 0
                st = lib.b2r_reduce_local_noalloc(x.data_ptr(), got.data_ptr(), n, DTYPES[dtype], op, stream, None)
                assert st == 0, lib.b2r_last_error_string().decode()
                torch.cuda.synchronize()
                mag = x.double().abs().sum().item()
                assert agree(got, want, dtype, op, n, mag), (dtype, op, n, got.item(), want.item())
    finally:
        lib.b2r_workspace_destroy(ws)


def test_noalloc_allocates_nothing_and_handles_edge_sizes(lib):
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    x = make_input(1 << 20, torch.float32, 3)
    out = torch.zeros((), dtype=torch.float32, device="cuda")
    assert lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), FLOAT32, SUM, stream, None) == 0  # set-up happened
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info()
    for _ in range(50):
        assert lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), FLOAT32, MAX, stream, None) == 0
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info()
    assert free_after == free_before, "device memory in use changed across calls that must not allocate"
    assert out.item() == x.max().item()
    # one element, and an empty input (identity)
    assert lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), 1, FLOAT32, SUM, stream, None) == 0
    torch.cuda.synchronize()
    assert out.item() == x[0].item()
    assert lib.b2r_reduce_local_noalloc(None, out.data_ptr(), 0, FLOAT32, MIN, stream, None) == 0
    torch.cuda.synchronize()
    assert out.item() == math.inf


def test_noalloc_rejects_geometries_beyond_the_static_scratch(lib):
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    x = make_input(1 << 22, torch.int32, 4)
    out = torch.zeros((), dtype=torch.int32, device="cuda")
    cfg = LaunchCfg()
    cfg.ctas = MAX_PARTIALS  # exactly at the limit: fine
    cfg.threads = 128
    assert lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), INT32, SUM, stream, C.byref(cfg)) == 0
    torch.cuda.synchronize()
    want = int(x.sum(dtype=torch.int64).item())
    assert out.item() == (want + 2**31) % 2**32 - 2**31
    cfg.ctas = MAX_PARTIALS + 1  # one more partial than the array holds
    st = lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), INT32, SUM, stream, C.byref(cfg))
    assert st == 1  # B2R_ERR_INVALID
    msg = lib.b2r_last_error_string().decode()
    assert "static scratch holds" in msg and str(MAX_PARTIALS) in msg
    cfg.ctas = 0
    cfg.variant = 3  # the TMA variant is not offered without a workspace
    assert lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), INT32, SUM, stream, C.byref(cfg)) == 1


def test_noalloc_never_uses_programmatic_launch_even_with_B2R_PDL_set(lib, monkeypatch):
    """Calls share one scratch array, so they must not overlap: with B2R_PDL=1 exported (the documented way to turn
    programmatic dependent launch on for the whole process) the no-alloc launches still run in plain stream order and agree
    bit for bit with the workspace entry point given the same static geometry."""
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    monkeypatch.setenv("B2R_PDL", "1")
    lib.b2r_refresh_env()
    ws = C.c_void_p()
    assert lib.b2r_workspace_create(C.byref(ws), 0) == 0
    try:
        same = LaunchCfg()
        # ldg128, plain order, static split, and an explicit grid below the 1024-partial limit: what the no-alloc path launches
        same.variant, same.pdl, same.dyn_pct, same.ctas, same.threads = 1, -1, -1, 592, 512
        asked = LaunchCfg()
        asked.ctas, asked.threads = 592, 512  # pdl left at 0 on purpose: "whatever B2R_PDL says" must still mean OFF here
        for dtype in DTYPES:
            outs = []
            for n, seed in (((1 << 23) + 5, 11), ((1 << 23) + 5, 12), (123_457, 13)):
                x = make_input(n, dtype, seed)
                want = torch.zeros((), dtype=dtype, device="cuda")
                got = torch.zeros((), dtype=dtype, device="cuda")
                assert lib.b2r_reduce_local_ex(x.data_ptr(), want.data_ptr(), n, DTYPES[dtype], SUM, ws, stream, C.byref(same)) == 0
                # two no-alloc calls back to back, no synchronisation in between
                assert lib.b2r_reduce_local_noalloc(x.data_ptr(), got.data_ptr(), n, DTYPES[dtype], SUM, stream, C.byref(asked)) == 0
                got2 = torch.zeros((), dtype=dtype, device="cuda")
                assert lib.b2r_reduce_local_noalloc(x.data_ptr(), got2.data_ptr(), n, DTYPES[dtype], SUM, stream, C.byref(asked)) == 0
                outs.append((x, want, got, got2))
            torch.cuda.synchronize()
            for x, want, got, got2 in outs:
                wb, gb, g2 = (t.cpu().reshape(1).view(torch.uint8) for t in (want, got, got2))
                assert torch.equal(wb, gb) and torch.equal(wb, g2), (dtype, want.item(), got.item(), got2.item())
    finally:
        lib.b2r_workspace_destroy(ws)
        monkeypatch.delenv("B2R_PDL")
        lib.b2r_refresh_env()


@pytest.mark.multigpu
def test_noalloc_on_a_second_device(lib):
    """Every device has its own copy of the static array and its own one-time set-up."""
    for dev in (1, 0, 1):
        torch.cuda.set_device(dev)
        stream = torch.cuda.current_stream().cuda_stream
        x = make_input((1 << 21) + 7, torch.int32, 20 + dev).to(f"cuda:{dev}")
        out = torch.zeros((), dtype=torch.int32, device=f"cuda:{dev}")
        st = lib.b2r_reduce_local_noalloc(x.data_ptr(), out.data_ptr(), x.numel(), INT32, SUM, stream, None)
        assert st == 0, lib.b2r_last_error_string().decode()
        torch.cuda.synchronize()
        want = int(x.sum(dtype=torch.int64).item())
        assert out.item() == (want + 2**31) % 2**32 - 2**31
    torch.cuda.set_device(0)
