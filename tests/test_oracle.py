"""CPU-only: the oracles and the generator (W1 of SURVEY §8). Oracle counterpart: reduction.cpp:214-249."""
import math

import pytest
import torch

import cuda_mpi_reductions_b200 as b2r

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]
SIZES = [0, 1, 2, 31, 32, 33, 255, 256, 257, 1023, 4097, 65537]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("n", SIZES)
def test_oracle_matches_torch(dtype, n):
    x = b2r.fill_host(n, dtype, seed=n + 7, mode="signed")
    if n == 0:
        assert b2r.reduce_oracle(x, "sum").item() == 0
        if dtype.is_floating_point:
            assert b2r.reduce_oracle(x, "min").item() == math.inf
            assert b2r.reduce_oracle(x, "max").item() == -math.inf
        else:
            assert b2r.reduce_oracle(x, "min").item() == 2**31 - 1
            assert b2r.reduce_oracle(x, "max").item() == -(2**31)
        return
    assert b2r.reduce_oracle(x, "min").item() == x.min().item()
    assert b2r.reduce_oracle(x, "max").item() == x.max().item()
    s = b2r.reduce_oracle(x, "sum")
    if dtype == torch.int32:
        want = x.to(torch.int64).sum().item()
        want = (want + 2**31) % 2**32 - 2**31
        assert s.item() == want
    elif dtype == torch.bfloat16:
        assert s.item() == x.double().sum().to(torch.bfloat16).item()
    else:
        ref = math.fsum(x.double().tolist())
        tol = 1e-15 * n if dtype == torch.float64 else 1e-7 * max(1.0, abs(ref))
        assert abs(s.item() - ref) <= tol + (0 if dtype == torch.float64 else abs(ref) * 6e-8)


def test_int_sum_wraps():
    x = torch.full((5,), 2**31 - 1, dtype=torch.int32)
    got = b2r.reduce_oracle(x, "sum").item()
    assert got == ((5 * (2**31 - 1) + 2**31) % 2**32) - 2**31


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.bfloat16])
def test_nan_and_inf_policy(dtype):
    x = torch.tensor([1.0, float("nan"), -3.0, float("inf")], dtype=dtype)
    assert math.isnan(b2r.reduce_oracle(x, "min").item())
    assert math.isnan(b2r.reduce_oracle(x, "max").item())
    y = torch.tensor([1.0, float("-inf"), -3.0, float("inf")], dtype=dtype)
    assert b2r.reduce_oracle(y, "min").item() == -math.inf
    assert b2r.reduce_oracle(y, "max").item() == math.inf
    assert math.isnan(b2r.reduce_oracle(y, "sum").item())


def test_kahan_beats_naive():
    # 1 + many tiny terms: a naive fp32 loop loses them all, the oracle must not
    x = torch.cat([torch.ones(1), torch.full((1 << 16,), 2.0**-25)]).float()
    got = b2r.reduce_oracle(x, "sum").item()
    assert abs(got - (1 + (1 << 16) * 2.0**-25)) < 1e-7


@pytest.mark.parametrize("dtype", DTYPES)
def test_fill_is_counter_based(dtype):
    a = b2r.fill_host(1000, dtype, seed=3, mode="full")
    b = b2r.fill_host(400, dtype, seed=3, first_index=600, mode="full")
    assert torch.equal(a[600:].view(torch.uint8), b.view(torch.uint8))
    c = b2r.fill_host(1000, dtype, seed=4, mode="full")
    assert not torch.equal(a.view(torch.uint8), c.view(torch.uint8))


def test_sdk_mode_is_small():
    x = b2r.fill_host(4096, torch.int32, mode="sdk")
    assert x.min() >= 0 and x.max() <= 255
    y = b2r.fill_host(4096, torch.bfloat16, mode="sdk")
    assert y.float().max() < 1.0 and y.float().min() >= 0.0


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("op", ["sum", "min", "max"])
def test_combine_oracle(dtype, op):
    a = b2r.fill_host(513, dtype, seed=1, mode="signed")
    c = b2r.fill_host(513, dtype, seed=2, mode="signed")
    acc = a.clone()
    b2r.combine_oracle(acc, c, op)
    if op == "sum":
        if dtype == torch.int32:
            want = ((a.long() + c.long() + 2**31) % 2**32 - 2**31).int()
        elif dtype == torch.bfloat16:
            want = (a.float() + c.float()).bfloat16()
        else:
            want = a + c
    else:
        want = torch.minimum(a, c) if op == "min" else torch.maximum(a, c)
    assert torch.equal(acc, want)
