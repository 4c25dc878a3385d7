"""CPU-only property tests (hypothesis): the pure host-side pieces under random inputs.

* the launch planner either rejects a hint set with B2RError or returns a geometry that satisfies every invariant the
  kernels rely on (the TMA tile bug of round 1 — tile sized for a thread count that was lowered afterwards — is exactly
  the kind of thing a fixed list of cases misses);
* the CPU oracles agree with an independent Python restatement for arbitrary contents, including NaN / inf / -0.0;
* the host collective algorithms agree with each other wherever the combination order cannot matter, for any rank count,
  root and length;
* the results aggregator (getAvgs.sh counterpart) is order-independent and matches exact decimal arithmetic."""
import math
import random
from decimal import Decimal

import torch
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200.parallel import host_reduce_threads
from cuda_mpi_reductions_b200.utils import results

DTYPES = [torch.int32, torch.float32, torch.float64, torch.bfloat16]
COMMON = dict(deadline=None, suppress_health_check=[HealthCheck.too_slow])


# ------------------------------------------------------------------------------------------------ planner
hint_sets = st.fixed_dictionaries(
    {},
    optional={
        "variant": st.sampled_from(["ldg128", "ldg256", "tma"]),
        "threads": st.sampled_from([32, 64, 96, 128, 160, 256, 288, 384, 512, 100, 544]),
        "ctas": st.sampled_from([1, 2, 3, 7, 148, 149, 296, 592, 1000, 4096, 4097]),
        "cluster": st.sampled_from([1, 2, 3, 4, 8, 16]),
        "stages": st.sampled_from([1, 2, 3, 4, 6, 8, 16, 32, 33]),
        "unroll": st.sampled_from([4, 8, 5]),
        "pdl": st.sampled_from([-1, 0, 1, 2]),
        "dyn_pct": st.sampled_from([-1, 0, 10, 50, 90, 200]),
    },
)


@settings(max_examples=400, **COMMON)
@given(n=st.one_of(st.integers(0, 1 << 14), st.integers(0, 1 << 40)), dtype=st.sampled_from(DTYPES), op=st.sampled_from(["sum", "min", "max"]),
       cfg=hint_sets)
def test_planner_output_always_satisfies_the_kernel_invariants(n, dtype, op, cfg):
    try:
        p = b2r.plan(n, dtype, op, cfg=cfg or None)
    except b2r.B2RError:
        return  # rejected hint sets are fine; silently producing a broken geometry is not
    assert p == b2r.plan(n, dtype, op, cfg=cfg or None)  # pure function
    assert p["threads"] % 32 == 0 and 32 <= p["threads"] <= 512
    assert p["cluster"] in (1, 2, 4, 8)
    assert 1 <= p["ctas"] <= 4096 and p["ctas"] % p["cluster"] == 0
    assert p["unroll"] in (4, 8)
    if p["variant"] == "ldg256":
        assert p["unroll"] == 4  # ptxas serialises eight 256-bit loads
    if p["variant"] == "tma":
        assert p["threads"] >= 64
        assert p["tile_bytes"] == (p["threads"] - 32) * 64  # every consumer thread owns 4 vectors of a tile
        assert 1 <= p["stages"] <= 32
        assert p["stages"] * p["tile_bytes"] + 16 * p["stages"] <= 220 * 1024
    # hints that were accepted are honoured
    for key in ("variant", "stages"):
        if key in cfg and (key != "stages" or p["variant"] == "tma"):
            assert p[key] == cfg[key], (key, cfg, p)
    assert p["pdl"] == {-1: 0, 0: 1, 1: 1, 2: 2}[cfg.get("pdl", 0)]  # 0 = default = the safe flavour
    assert 0 <= p["dyn_pct"] <= 90
    if p["dyn_pct"]:
        assert p["variant"] != "tma" and p["cluster"] == 1 and p["pdl"] != 2 and p["ctas"] >= 2
    if "ctas" in cfg:
        assert p["ctas"] in (cfg["ctas"], cfg["ctas"] - cfg["ctas"] % p["cluster"])
        if "threads" in cfg:
            assert p["threads"] == cfg["threads"]
    if n * torch.empty(0, dtype=dtype).element_size() <= 64 * 1024 and "ctas" not in cfg:
        assert p["ctas"] == 1 and p["cluster"] == 1  # no ticket stage for inputs a single CTA streams in a few microseconds


# ------------------------------------------------------------------------------------------------ oracles
def py_reduce(vals, op, is_int):
    if op == "sum":
        if is_int:
            s = sum(vals) & 0xFFFFFFFF
            return s - (1 << 32) if s >= (1 << 31) else s
        return math.fsum(vals)
    if any(isinstance(v, float) and math.isnan(v) for v in vals):
        return math.nan
    return min(vals) if op == "min" else max(vals)


special_floats = st.sampled_from([0.0, -0.0, 1.0, -1.0, math.inf, -math.inf, math.nan, 1e30, -1e30, 1e-30])


@settings(max_examples=150, **COMMON)
@given(data=st.data(), dtype=st.sampled_from(DTYPES), op=st.sampled_from(["sum", "min", "max"]))
def test_reduce_oracle_against_python(data, dtype, op):
    if dtype == torch.int32:
        vals = data.draw(st.lists(st.integers(-(1 << 31), (1 << 31) - 1), min_size=1, max_size=300))
        x = torch.tensor(vals, dtype=torch.int32)
        assert b2r.reduce_oracle(x, op).item() == py_reduce(vals, op, True)
        return
    elems = st.one_of(st.floats(-1e6, 1e6, width=32), special_floats) if op != "sum" else st.floats(-1e6, 1e6, width=32)
    vals = data.draw(st.lists(elems, min_size=1, max_size=300))
    x = torch.tensor(vals, dtype=torch.float64).to(dtype)
    exact = [float(v) for v in x.double().tolist()]  # what the oracle actually sees after the cast
    got = b2r.reduce_oracle(x, op).double().item()
    want = py_reduce(exact, op, False)
    if math.isnan(want):
        assert math.isnan(got)
    elif op == "sum":
        # compensated summation, rounded once to the output type
        ulp = {torch.float32: 2.0**-23, torch.float64: 2.0**-52, torch.bfloat16: 2.0**-7}[dtype]
        assert abs(got - want) <= 2 * ulp * max(abs(want), sum(abs(v) for v in exact) * 2.0**-20, 1e-30)
    else:
        assert got == want and math.copysign(1, got) in (math.copysign(1, want), -math.copysign(1, want))


# ------------------------------------------------------------------------------------------------ host collectives
@settings(max_examples=60, **COMMON)
@given(nranks=st.integers(1, 9), count=st.integers(1, 700), dtype=st.sampled_from(DTYPES), op=st.sampled_from(["sum", "min", "max"]),
       seed=st.integers(0, 1 << 30), root_pick=st.integers(0, 1 << 10))
def test_host_algorithms_agree_where_order_cannot_matter(nranks, count, dtype, op, seed, root_pick):
    xs = [b2r.fill_host(count, dtype, seed=seed + 977 * r, mode="signed") for r in range(nranks)]
    root = root_pick % nranks
    order_free = dtype == torch.int32 or op != "sum"
    ref = host_reduce_threads(xs, op, root, "oneshot")
    for algo in ("twoshot", "tree", "ring", "push", "ll"):
        got = host_reduce_threads(xs, op, root, algo)
        if order_free or algo in ("twoshot", "push", "ll"):  # these fold in rank order, exactly like one-shot
            assert torch.equal(got.view(torch.uint8), ref.view(torch.uint8)), (algo, nranks, count, dtype, op, root)
        else:
            tol = {torch.float32: 1e-5, torch.float64: 1e-13, torch.bfloat16: 0.1}[dtype] * nranks
            assert torch.allclose(got.double(), ref.double(), rtol=tol, atol=tol * 50), (algo, nranks, count, dtype, op, root)
    # allreduce: every rank's row equals the reduce result
    rows = host_reduce_threads(xs, op, -1, "twoshot")
    for r in range(nranks):
        assert torch.equal(rows[r].view(torch.uint8), ref.view(torch.uint8))


# ------------------------------------------------------------------------------------------------ results aggregator
@settings(max_examples=80, **COMMON)
@given(rows=st.lists(st.tuples(st.sampled_from(["INT", "DOUBLE"]), st.sampled_from(["MAX", "MIN", "SUM"]), st.sampled_from([2, 4, 8, 64, 1024]),
                               st.decimals(min_value=Decimal("0.001"), max_value=Decimal("99999"), places=6)), min_size=1, max_size=60),
       seed=st.integers(0, 1000))
def test_aggregator_is_order_independent_and_exact(tmp_path_factory, rows, seed):
    d = tmp_path_factory.mktemp("agg")
    lines = ["# DATATYPE OP NODES GB/sec"] + [f"{t} {o} {n} {v}" for t, o, n, v in rows]
    a = d / "a.txt"
    a.write_text("\n".join(lines) + "\n")
    shuffled = lines[1:]
    random.Random(seed).shuffle(shuffled)
    b = d / "b.txt"
    b.write_text("\n".join(["garbage line that is not a row"] + shuffled) + "\n")
    def read_back(collected, out):
        got = {}
        for (t, o), path in results.aggregate(str(collected), str(out)).items():
            for line in open(path):
                f = line.split()
                if len(f) == 4:
                    assert (f[0], f[1]) == (t, o)
                    got[(t, o, int(f[2]))] = f[3]
        return got

    ra = read_back(a, d / "out_a")
    rb = read_back(b, d / "out_b")
    assert ra == rb
    groups = {}
    for t, o, n, v in rows:
        groups.setdefault((t, o, n), []).append(v)
    assert set(ra) == set(groups)
    for key, vals in groups.items():
        mean = sum(vals) / len(vals)
        got = Decimal(ra[key])
        assert abs(got - mean) < Decimal("0.00001") and got <= mean  # truncated to 5 digits, never rounded up
