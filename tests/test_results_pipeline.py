"""CPU-only: result aggregation and plot-script generation (reference: mpi/getAvgs.sh:3-14, mpi/makePlots.gp:1-39) and
the mini-MPI plumbing configuration (BASELINE.json config 1: custom tree vs MPI_Reduce on CPU ranks)."""
import os
import subprocess

import pytest

from cuda_mpi_reductions_b200.models import MPI_REDUCE, SDK_REDUCTION
from cuda_mpi_reductions_b200.utils import results

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_aggregate_matches_reference_layout(tmp_path):
    collected = tmp_path / "collected.txt"
    collected.write_text(
        "# DATATYPE OP NODES GB/sec\n"
        "INT MAX 64      8.260\nINT MAX 64      8.520\nINT SUM 256     38.0\nINT SUM 64      9.126\nINT SUM 64      9.238\n"
        "DOUBLE MIN 1024  88.7626\nnoise line\nINT SUM 256     39.29681\n")
    out = results.aggregate(str(collected), str(tmp_path / "results"))
    text = open(out[("INT", "SUM")]).read()
    assert text == "\nINT SUM 64 9.18200\nINT SUM 256 38.64840\n"  # blank first line, ascending nodes, 5 decimals truncated
    assert open(out[("INT", "MAX")]).read() == "\nINT MAX 64 8.39000\n"
    assert ("DOUBLE", "MIN") in out


def test_reference_results_are_reproduced_from_its_collected_file(tmp_path):
    ref = "/root/reference/mpi/collected.txt"
    if not os.path.exists(ref):
        pytest.skip("reference tree not mounted")
    out = results.aggregate(ref, str(tmp_path))
    got = open(out[("INT", "SUM")]).read().split("\n")
    want = open("/root/reference/mpi/results/INT_SUM.txt").read().split("\n")
    assert sorted(l for l in got if l) == sorted(l for l in want if l)  # the reference lists nodes in `sort -n | uniq` order too


def test_gnuplot_script(tmp_path):
    (tmp_path / "INT_SUM.txt").write_text("\nINT SUM 2 100.0\n")
    path = results.write_gnuplot(str(tmp_path), str(tmp_path / "p.gp"), {("INT", "SUM"): 7000.0, ("INT", "MIN"): 6900.0, ("INT", "MAX"): 6950.0})
    gp = open(path).read()
    assert "int_reduce.svg" in gp and "INT_SUM.txt" in gp and "one GPU SUM" in gp and "double_reduce.svg" in gp
    assert f"{7000.0 / 1.073741824:.3f}" in gp  # single-GPU GB/s converted to the GiB/s of the curves


def test_workload_constants_match_the_reference():
    assert SDK_REDUCTION.elements == 1 << 24 and SDK_REDUCTION.nbytes("int") == 64 << 20
    assert MPI_REDUCE.count_per_rank("int", 8) == (512 << 20) // 8 and MPI_REDUCE.count_per_rank("double", 1024) == (256 << 20) // 1024
    assert list(MPI_REDUCE.sweep_bytes())[0] == 1024 and list(MPI_REDUCE.sweep_bytes())[-1] == 1 << 32


@pytest.mark.parametrize("np_", [2, 3, 4])
@pytest.mark.parametrize("algo", ["twoshot", "oneshot", "tree"])
def test_minimpi_custom_tree_vs_library_reduce(np_, algo):
    exe, run = os.path.join(REPO, "bin", "tree_reduce"), os.path.join(REPO, "bin", "b2r_mpirun")
    if not (os.path.exists(exe) and os.path.exists(run)):
        subprocess.run(["make", "-C", os.path.join(REPO, "minimpi")], check=True, capture_output=True)
    env = dict(os.environ, MINIMPI_ALGO=algo)
    res = subprocess.run([run, "-np", str(np_), "--window-mb", "1", exe, "600000", "1"], capture_output=True, text=True, timeout=120, env=env)
    assert res.returncode == 0 and "PASSED" in res.stdout, res.stdout[-500:] + res.stderr[-500:]
    rows = list(results.parse_rows(res.stdout.splitlines()))
    assert {r[0] for r in rows} == {"TREE_INT", "LIB_INT", "TREE_DOUBLE", "LIB_DOUBLE"} and all(r[2] == np_ for r in rows)


def test_minimpi_aborts_cleanly_when_not_launched():
    exe = os.path.join(REPO, "bin", "tree_reduce")
    if not os.path.exists(exe):
        pytest.skip("not built")
    env = {k: v for k, v in os.environ.items() if not k.startswith("MINIMPI_")}
    res = subprocess.run([exe], capture_output=True, text=True, timeout=30, env=env)
    assert res.returncode != 0 and "b2r_mpirun" in res.stderr


def test_compare_sweeps_tool(tmp_path):
    """scripts/compare_sweeps.py: parses both tools' sweep rows, keeps the fastest duplicate, ranks ours against NCCL and
    reports the size ranges each algorithm wins."""
    import subprocess
    import sys

    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ours = tmp_path / "ours.txt"
    ours.write_text(
        "# coll      dtype  op   algo              bytes         us algbw GB/s busbw GB/s\n"
        "  reduce    FLOAT  SUM  ll                 1024       6.00       0.17       0.17\n"
        "  reduce    FLOAT  SUM  nvls               1024      12.50       0.08       0.08\n"
        "  reduce    FLOAT  SUM  nvls            1048576      17.80      58.90      58.90\n"
        "  reduce    FLOAT  SUM  ring            1048576      30.00      34.95      34.95\n"
        "  reduce    FLOAT  SUM  nvls          268435456     522.00     514.24     514.24\n"
        "  reduce    FLOAT  SUM  ring          268435456     430.00     624.27     624.27\n"
        "  reduce    FLOAT  SUM  ring          268435456     440.00     610.08     610.08\n"
        "PASSED\n")
    nccl = tmp_path / "nccl.txt"
    nccl.write_text("  reduce    FLOAT  SUM  nccl               1024       8.30       0.12       0.12\n"
                    "  reduce    FLOAT  SUM  nccl          268435456     445.70     602.28     602.28\n")
    md = tmp_path / "t.md"
    res = subprocess.run([sys.executable, os.path.join(repo, "scripts/compare_sweeps.py"), str(ours), str(nccl), "--md", str(md)],
                         capture_output=True, text=True)
    assert res.returncode == 0, res.stderr
    out = res.stdout
    assert "1 KiB: ll (>= 1.38x NCCL)" in out and "1 MiB: nvls" in out and "256 MiB: ring (>= 1.04x NCCL)" in out
    assert md.read_text().count("\n") == 5  # header, rule, three sizes


def test_svg_figures_are_written_without_gnuplot(tmp_path):
    """The reference's two figures (mpi/makePlots.gp:21-39) as self-contained SVG: three rank counts per curve like
    mpi/results/INT_SUM.txt:2-4, implementation families side by side, single-GPU throughput as dashed lines."""
    import xml.dom.minidom

    from cuda_mpi_reductions_b200.utils import results as R

    for sub in ("", "nccl", "tree"):
        d = tmp_path / sub
        d.mkdir(exist_ok=True)
        for dt in ("INT", "DOUBLE"):
            for k, op in enumerate(R.OPS):
                (d / f"{dt}_{op}.txt").write_text("\n" + "".join(f"{dt} {op} {p} {100.0 * p * (k + 1) + len(sub):.5f}\n" for p in (2, 4, 8)))
    written = R.write_figures(str(tmp_path), {("INT", "SUM"): 6000.0, ("DOUBLE", "MAX"): 5000.0})
    assert [os.path.basename(w) for w in written] == ["int.svg", "double.svg"]
    for w in written:
        doc = xml.dom.minidom.parse(w)
        assert len(doc.getElementsByTagName("path")) == 9          # 3 families x 3 ops
        assert len(doc.getElementsByTagName("circle")) >= 27       # three rank counts per curve (+ legend markers)
    assert "one GPU SUM" in open(written[0]).read() and "stroke-dasharray" in open(written[0]).read()
    with pytest.raises(ValueError):
        R.write_svg({}, str(tmp_path / "empty.svg"), "nothing")


def test_sweep_figures_from_harness_csv(tmp_path):
    """BASELINE config 5 as figures: rows of `bench_collective --sweep --csv` / `nccl_bench --sweep --csv` -> one SVG per (GPU count,
    collective) with a b2r and an NCCL line per (dtype, op); re-measured plan rows replace b2r rows up to a size limit."""
    import xml.dom.minidom

    from cuda_mpi_reductions_b200.utils import results as R

    sizes = [1 << 10, 1 << 20, 1 << 30]
    base = tmp_path / "sweep.csv"
    with open(base, "w") as f:
        for impl, scale in (("b2r", 2.0), ("nccl", 1.0)):
            for coll in ("reduce", "allreduce"):
                for dt in ("INT", "BF16"):
                    for b in sizes:
                        us = b / 1e3 / scale + 10
                        f.write(f"{impl},{coll},{dt},SUM,{'auto' if impl == 'b2r' else 'nccl'},8,{b},{us:.3f},{b / us * 1e-3:.3f},{b / us * 1e-3:.3f}\n")
    plan = tmp_path / "plan.csv"
    plan.write_text(f"b2r,reduce,INT,SUM,posted,8,{1 << 20},5.000,209.715,209.715,R.posted,\n"
                    f"b2r,reduce,INT,SUM,posted,8,{1 << 30},5.000,1.000,1.000,R.posted,\n"      # beyond the limit: ignored
                    f"b2r,reduce,FLOAT,MIN,posted,8,{1 << 20},5.000,1.000,1.000,R.posted,\n"     # no such series in the sweep: ignored
                    f"b2r,reduce,INT,SUM,nvls,8,{1 << 20},9.000,1.000,1.000,Q.base,\n")          # another label: ignored
    rows = R.read_sweep_csv([str(base)], [(str(plan), "R.posted", str(8 << 20))])
    assert rows[(8, "reduce", "b2r", "INT", "SUM")][1 << 20] == (5.0, 209.715)
    assert rows[(8, "reduce", "b2r", "INT", "SUM")][1 << 30][0] != 5.0
    assert (8, "reduce", "b2r", "FLOAT", "MIN") not in rows
    written = R.write_sweep_figures(rows, str(tmp_path / "figs"))
    assert sorted(os.path.basename(w) for w in written) == ["sweep_allreduce_8gpu.svg", "sweep_reduce_8gpu.svg"]
    for w in written:
        doc = xml.dom.minidom.parse(w)
        assert len(doc.getElementsByTagName("path")) == 4  # b2r + NCCL for two dtypes
        text = open(w).read()
        assert ">1 KiB<" in text and ">1 MiB<" in text and ">1 GiB<" in text and "NCCL BF16 SUM" in text


def test_native_harness_defaults_are_the_reference_workloads():
    """models/workloads.py is the single table of the reference's workload constants; the C++ harnesses cannot import it, so
    their flag defaults are checked against it here (and bench.py loads it by path)."""
    import re

    src_c = open(os.path.join(REPO, "bench", "bench_collective.cu")).read()
    src_r = open(os.path.join(REPO, "bench", "bench_reduce.cu")).read()
    total_mib = int(re.search(r'i64\("total-mib", (\d+)\)', src_c).group(1))
    retries = int(re.search(r'i64\("retries", (\d+)\)', src_c).group(1))
    assert total_mib << 20 == MPI_REDUCE.total_ints * 4 == MPI_REDUCE.total_doubles * 8
    assert retries == MPI_REDUCE.retries
    n_default = re.search(r'i64\("n", ([^)]+)\)', src_r).group(1)
    assert eval(n_default.replace("ll", "")) == SDK_REDUCTION.elements  # noqa: S307 (a literal like 1 << 24)
    bench_py = open(os.path.join(REPO, "bench.py")).read()
    assert "SDK_REDUCTION.elements" in bench_py and "SDK_REDUCTION.published_gbs" in bench_py
