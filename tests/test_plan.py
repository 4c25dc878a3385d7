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


def test_pdl_plans_one_cta_per_sm():
    p = b2r.plan(1 << 24, torch.int32, cfg={"pdl": 1})
    assert p["pdl"] == 1 and p["ctas"] == 148
    assert b2r.plan(1 << 24, torch.int32)["pdl"] == 0


def test_tma_tile_matches_final_thread_count():
    for n in (256, 1024, 2048, 4096, 1 << 16, 1 << 24):
        for cfg in ({"variant": "tma"}, {"variant": "tma", "pdl": 1}, {"variant": "tma", "threads": 512}):
            p = b2r.plan(n, torch.float64, cfg=cfg)
            assert p["tile_bytes"] == (p["threads"] - 32) * 64, (n, cfg, p)
