"""CPU-only: the Python surface behaves sensibly without a GPU — helpers work, GPU entry points fail loudly instead of
silently falling back (the driver records which native libraries the GPU tests really loaded)."""
import ctypes as C
import os

import pytest
import torch

import cuda_mpi_reductions_b200 as b2r
from cuda_mpi_reductions_b200 import _lib
from cuda_mpi_reductions_b200.utils import ClockSampler, rotating_copies


def test_native_library_is_the_in_tree_build():
    assert os.path.exists(_lib.LIB_PATH) and _lib.LIB_PATH.startswith(_lib.REPO_ROOT)
    assert _lib.LIB.b2r_version() >= 1
    for name in ("b2r_reduce_local", "b2r_reduce_local_noalloc", "b2r_allreduce", "b2r_reduce", "b2r_reduce_scalar_sharded_ex",
                 "b2r_comm_init_rank", "b2r_comm_init_all", "b2r_comm_init_loopback", "b2r_host_comm_init", "b2r_bootstrap_selftest"):
        assert hasattr(_lib.LIB, name), name


def test_cpu_tensors_are_rejected_by_gpu_ops():
    x = torch.ones(16)
    with pytest.raises(ValueError):
        b2r.reduce(x, "sum")
    with pytest.raises(TypeError):
        b2r.reduce_oracle(torch.ones(4, dtype=torch.float16), "sum")
    with pytest.raises(ValueError):
        b2r.reduce_oracle(x, "mean")


def test_gpu_entry_points_fail_loudly_without_a_device():
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = C.c_void_p()
    st = _lib.LIB.b2r_workspace_create(C.byref(h), 0)
    assert st == _lib.ERR_CUDA and b"b2r_workspace_create" in _lib.LIB.b2r_last_error_string() or st == _lib.ERR_CUDA
    arr = (C.c_void_p * 2)()
    assert _lib.LIB.b2r_comm_init_loopback(arr, 2, 0, None) != _lib.OK
    assert _lib.LIB.b2r_comm_init_all(arr, 2, None, None) != _lib.OK


def test_status_strings_and_launch_config():
    assert _lib.LIB.b2r_status_string(_lib.ERR_TIMEOUT).decode() == "cross-GPU wait timed out"
    cfg = b2r.launch_config(variant="tma", threads=288, pdl=1)
    assert (cfg.variant, cfg.threads, cfg.pdl) == (_lib.VARIANT_TMA, 288, 1)
    with pytest.raises(TypeError):
        b2r.launch_config(blocks=3)


def test_rotating_copies_and_clock_sampler_fallback():
    assert rotating_copies(64 << 20) == 8 and rotating_copies(1 << 30) == 2 and rotating_copies(1024) == 64 and rotating_copies(0) == 1
    s = ClockSampler(0).start()
    s.stop()
    summ = s.summary(0.0, 1e12)
    assert set(summ) >= {"sm_mhz", "sm_max_mhz", "reasons", "samples", "window", "backend"}


def test_numa_binding_helpers(tmp_path, monkeypatch):
    """utils/numa.py on a fake sysfs tree: cpulist parsing, node lookup, and every 'leave it alone' exit."""
    import os

    from cuda_mpi_reductions_b200.utils import numa

    assert numa.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert numa.parse_cpulist("") == set()
    allowed = sorted(os.sched_getaffinity(0))
    root = tmp_path / "sys"
    (root / "bus/pci/devices/0000:1b:00.0").mkdir(parents=True)
    (root / "bus/pci/devices/0000:1b:00.0/numa_node").write_text("1\n")
    (root / "bus/pci/devices/0000:9c:00.0").mkdir(parents=True)
    (root / "bus/pci/devices/0000:9c:00.0/numa_node").write_text("-1\n")
    (root / "devices/system/node/node1").mkdir(parents=True)
    (root / "devices/system/node/node1/cpulist").write_text(f"{allowed[0]}\n")
    assert numa.gpu_numa_node("0000:1b:00.0", str(root)) == 1
    assert numa.gpu_numa_node("0000:9c:00.0", str(root)) is None  # the platform does not say
    assert numa.gpu_numa_node("0000:aa:00.0", str(root)) is None  # no such device
    info = numa.bind_to_gpu_node(0, str(root), pci_addr="0000:1b:00.0", apply=False)  # dry run: the test process keeps its CPUs
    assert info["node"] == 1 and info["cpus"] == 1 and info["bound"]
    assert not numa.bind_to_gpu_node(0, str(root), pci_addr="0000:9c:00.0", apply=False)["bound"]
    (root / "devices/system/node/node1/cpulist").write_text("100000\n")  # a CPU outside this process's cpuset
    assert not numa.bind_to_gpu_node(0, str(root), pci_addr="0000:1b:00.0", apply=False)["bound"]
    monkeypatch.setenv("B2R_NUMA_BIND", "0")
    assert numa.bind_to_gpu_node(0, str(root), pci_addr="0000:1b:00.0")["why"].startswith("disabled")
    monkeypatch.delenv("B2R_NUMA_BIND")
    assert numa.bind_to_gpu_node(0, apply=False)["bound"] in (True, False)  # real sysfs, with or without a GPU: must not raise
    assert sorted(os.sched_getaffinity(0)) == allowed


def test_cached_environment_lookups(tmp_path, monkeypatch):
    """src/util/env.h: B2R_* variables are read once per call site and again after b2r_refresh_env() — first as a C++ unit
    test of the header, then through the library (the introspection calls refresh on their own)."""
    import shutil
    import subprocess

    import torch

    import cuda_mpi_reductions_b200 as b2r

    repo = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    gxx = shutil.which("g++")
    if gxx:
        exe = tmp_path / "env_cache_test"
        subprocess.run([gxx, "-std=c++17", "-O1", "-I", os.path.join(repo, "src"), os.path.join(repo, "tests/cpp/env_cache_test.cpp"), "-o", str(exe)],
                       check=True)
        res = subprocess.run([str(exe)], capture_output=True, text=True)
        assert res.returncode == 0 and res.stdout.strip() == "OK", res.stderr
    monkeypatch.setenv("B2R_THREADS", "96")
    assert b2r.plan(1 << 24, torch.int32)["threads"] == 96  # plan() looks at the environment as it is now
    monkeypatch.delenv("B2R_THREADS")
    b2r.refresh_env()
    assert b2r.plan(1 << 24, torch.int32)["threads"] != 96


def test_prepared_scalar_sharded_call_passes_the_same_arguments(monkeypatch):
    """Communicator.prepare_reduce_scalar_sharded converts its arguments once; the native call must receive exactly what
    the regular method passes (checked with the native entry point and the stream lookup replaced by recorders)."""
    import ctypes as C

    import torch

    import cuda_mpi_reductions_b200 as b2r
    from cuda_mpi_reductions_b200.parallel import comm as cm

    calls = []

    def record(*a):
        vals = []
        for v in a:
            if v is None or isinstance(v, int):
                vals.append(v)
            elif hasattr(v, "value"):
                vals.append(v.value)
            else:  # byref(cfg): compare the pointed-to launch configuration
                vals.append(bytes(C.string_at(C.addressof(v._obj), C.sizeof(v._obj))))
        calls.append(tuple(0 if v is None else v for v in vals))
        return 0

    monkeypatch.setattr(cm, "_SCALAR_EX", record)
    monkeypatch.setattr(cm, "_raw_stream", lambda idx: 0xABC0)
    c = object.__new__(cm.Communicator)
    c.device, c.device_index, c._h, c.rank = torch.device("cpu"), 0, C.c_void_p(0x1234), 0
    xs = [torch.zeros(1000, dtype=torch.float64) for _ in range(3)]
    out = torch.zeros((), dtype=torch.float64)
    cfg = b2r.launch_config(pdl=1)
    for x in xs:
        c.reduce_scalar_sharded(x, "max", root=0, out=out, cfg=cfg)
    regular, calls[:] = list(calls), []
    start, out2 = c.prepare_reduce_scalar_sharded(xs, "max", root=0, out=out, cfg=cfg)
    assert out2 is out
    for i in range(5):
        start(i)
    assert calls[:3] == regular and calls[3:] == regular[:2]  # wraps around the buffer list
    with pytest.raises(ValueError):
        c.prepare_reduce_scalar_sharded([xs[0], torch.zeros(7, dtype=torch.float64)], "max")
