"""CPU-only: the launch planner (dispatcher) is a pure function of (n, dtype, op, SM count, hints).
Reference counterpart: getNumBlocksAndThreads (reduction.cpp:272-291)."""
import pytest
import torch

import cuda_mpi_reductions_b200 as b2r


def test_small_inputs_use_one_cta():
    p = b2r.plan(100, torch.float32)
    assert p["ctas"] == 1 and p["cluster"] == 1 and p["threads"] <= 128
    assert b2r.plan(0, torch.int32)["ctas"] == 1


def test_large_inputs_fill_the_machine():
    p = b2r.plan(1 << 24, torch.int32)
    assert p["ctas"] % 148 == 0 and p["ctas"] >= 148
    assert p["threads"] % 32 == 0


def test_grid_grows_with_size():
    sizes = [1 << k for k in range(10, 25, 2)]
    ctas = [b2r.plan(n, torch.float64)["ctas"] for n in sizes]
    assert ctas == sorted(ctas)
    assert b2r.plan(1 << 30, torch.float64)["ctas"] % 148 == 0


@pytest.mark.parametrize("cluster", [2, 4, 8])
def test_cluster_divides_grid(cluster):
    p = b2r.plan(1 << 26, torch.float32, cfg={"cluster": cluster})
    assert p["cluster"] == cluster and p["ctas"] % cluster == 0


def test_hints_are_respected_and_validated():
    p = b2r.plan(1 << 20, torch.bfloat16, cfg={"variant": "tma", "threads": 288, "stages": 6, "ctas": 148})
    assert p["variant"] == "tma" and p["threads"] == 288 and p["stages"] == 6 and p["ctas"] == 148
    assert p["tile_bytes"] == (288 - 32) * 64
    with pytest.raises(b2r.B2RError):
        b2r.plan(1 << 20, torch.float32, cfg={"threads": 100})
    with pytest.raises(b2r.B2RError):
        b2r.plan(1 << 20, torch.float32, cfg={"cluster": 3})
    with pytest.raises(b2r.B2RError):
        b2r.plan(1 << 20, torch.float32, cfg={"variant": "tma", "threads": 512, "stages": 32})


def test_env_override(monkeypatch):
    monkeypatch.setenv("B2R_VARIANT", "2")
    assert b2r.plan(1 << 24, torch.float32)["variant"] == "ldg256"


def test_pdl_flavours(monkeypatch):
    """pdl=2 (the caller declares the input complete) streams early and wants one CTA per SM; pdl=1 waits before its first
    load, overlaps nothing but launch latency, and therefore keeps the plain-order geometry. The environment can only ever
    ask for the safe flavour (an env switch must not turn on early reads at every call site)."""
    p = b2r.plan(1 << 24, torch.int32, cfg={"pdl": b2r.PDL_INPUT_READY})
    assert p["pdl"] == 2 and p["ctas"] == 148 and p["dyn_pct"] == 0
    default = b2r.plan(1 << 24, torch.int32)
    assert default["pdl"] == 1, "the safe flavour is the default"
    safe = b2r.plan(1 << 24, torch.int32, cfg={"pdl": 1})
    assert safe == default
    assert b2r.plan(1 << 24, torch.int32, cfg={"pdl": -1})["pdl"] == 0
    for v in ("1", "2", "7"):
        monkeypatch.setenv("B2R_PDL", v)
        assert b2r.plan(1 << 24, torch.int32)["pdl"] == 1, v
        assert b2r.plan(1 << 24, torch.int32, cfg={"pdl": -1})["pdl"] == 0, "a negative hint switches PDL off whatever the environment says"
    monkeypatch.setenv("B2R_PDL", "0")
    off = b2r.plan(1 << 24, torch.int32)
    assert off["pdl"] == 0 and off["ctas"] == default["ctas"] and off["variant"] == default["variant"]
    assert b2r.plan(1 << 24, torch.int32, cfg={"pdl": 1})["pdl"] == 1


def test_dynamic_tail_is_planned_for_mid_sizes_only(monkeypatch):
    assert b2r.plan(1 << 24, torch.int32)["dyn_pct"] == 25          # 64 MiB, plain order, ldg128
    assert b2r.plan(1 << 16, torch.int32)["dyn_pct"] == 0           # small: nothing to balance
    assert b2r.plan(1 << 28, torch.float32)["dyn_pct"] == 0         # 1 GiB: TMA ring, static split
    assert b2r.plan(1 << 24, torch.int32, cfg={"dyn_pct": -1})["dyn_pct"] == 0
    assert b2r.plan(1 << 24, torch.int32, cfg={"dyn_pct": 40})["dyn_pct"] == 40
    assert b2r.plan(1 << 24, torch.int32, cfg={"cluster": 2})["dyn_pct"] == 0
    assert b2r.plan(1 << 24, torch.int32, cfg={"variant": "tma"})["dyn_pct"] == 0
    monkeypatch.setenv("B2R_DYN_PCT", "10")
    assert b2r.plan(1 << 24, torch.int32)["dyn_pct"] == 10


def test_tma_tile_matches_final_thread_count():
    for n in (256, 1024, 2048, 4096, 1 << 16, 1 << 24):
        for cfg in ({"variant": "tma"}, {"variant": "tma", "pdl": 2}, {"variant": "tma", "threads": 512}):
            p = b2r.plan(n, torch.float64, cfg=cfg)
            assert p["tile_bytes"] == (p["threads"] - 32) * 64, (n, cfg, p)
