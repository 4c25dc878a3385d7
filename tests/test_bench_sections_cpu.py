"""CPU-only: the control flow of bench.py's `collectives` section (utils/bench_sections.py) with stand-ins for the communicator,
torch.distributed and the CUDA events — the section itself only runs at N > 1 on GPUs, where a slip in its Python would cost the
driver's record of every row. Checked: the rows it emits (the six >= 64 MiB cases plus the mid-size reduce row, marked
l2_resident), their self-checks, and that a failure of the mid-size row leaves the others standing."""
import types

import torch

from cuda_mpi_reductions_b200.utils import bench_sections as S

WORLD = 4


class FakeEvent:
    t = 0.0

    def __init__(self, enable_timing=True):
        self.at = 0.0

    def record(self):
        FakeEvent.t += 1.0
        self.at = FakeEvent.t

    def elapsed_time(self, other):
        return max(other.at - self.at, 1e-3)


class FakeDist:
    class ReduceOp:
        SUM, MIN, MAX = "sum", "min", "max"

    @staticmethod
    def _total(t):  # what the sum over WORLD ranks of _pattern(n, rank) is, given rank 0's pattern
        return t * WORLD + WORLD * (WORLD - 1) // 2

    def reduce(self, t, dst=0, op=None):
        t.copy_(self._total(t))

    def all_reduce(self, t, op=None):
        if op == "sum":
            t.copy_(self._total(t))
        # MIN / MAX over identical ranks: unchanged


class FakeComm:
    def __init__(self, fail_on_bytes=None):
        self.fail_on_bytes = fail_on_bytes
        self.calls = []

    def symm_empty(self, n, dtype):
        return torch.empty(n, dtype=dtype)

    def _do(self, x, out):
        if self.fail_on_bytes is not None and x.numel() * 4 == self.fail_on_bytes:
            raise RuntimeError("injected")
        out.copy_(FakeDist._total(x))

    def reduce(self, x, op, root=0, out=None):
        self.calls.append(("reduce", x.numel()))
        self._do(x, out)

    def allreduce(self, x, op, out=None):
        self.calls.append(("allreduce", x.numel()))
        self._do(x, out)

    def check(self):
        pass

    def barrier(self):
        pass

    def select_algo(self, nbytes, dtype, op, allreduce=True):
        return "nvls" if nbytes > (8 << 20) or allreduce else "posted"


def run(monkeypatch, comm):
    monkeypatch.setattr(torch.cuda, "Event", FakeEvent)
    monkeypatch.setattr(torch.cuda, "synchronize", lambda *a, **k: None)
    return S.collectives_section(comm, FakeDist(), torch.device("cpu"), 0, WORLD, sizes=(1 << 16, 8 << 20), iters_small=2, iters_big=1)


def test_collectives_section_rows(monkeypatch):
    out = run(monkeypatch, FakeComm())
    rows = out["rows"]
    assert [(r["coll"], r["dtype"], r["bytes"]) for r in rows] == [
        ("reduce", "float32", 1 << 16), ("reduce", "float32", 8 << 20), ("allreduce", "float32", 1 << 16), ("allreduce", "float32", 8 << 20),
        ("reduce", "int32", 1 << 16), ("allreduce", "int32", 1 << 16), ("reduce", "float32", 4 << 20)]
    assert all(r["b2r_check"] == "PASSED" and r["nccl_check"] == "PASSED" and r["ratio"] > 0 for r in rows)
    assert rows[-1]["l2_resident"] is True and rows[-1]["algo"] == "posted" and "l2_resident" not in rows[0]
    assert out["mid_size_row_error"] is None and out["timed_launches_per_gpu"] == 7 * 2


def test_a_failing_mid_size_row_leaves_the_others(monkeypatch):
    out = run(monkeypatch, FakeComm(fail_on_bytes=4 << 20))
    assert len(out["rows"]) == 6 and "injected" in out["mid_size_row_error"]
