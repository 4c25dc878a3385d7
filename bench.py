#!/usr/bin/env python
"""bench.py — headline benchmark of b200reduce (driver contract: see the task statement).

Metric  : reduction throughput in decimal GB/s, the number the reference prints as
          "Reduction, Throughput = %.4f GB/s" (cuda/C/src/reduction/reduction.cpp:744-745) and plots as the CUDA lines
          of its figures (mpi/makePlots.gp:17-19): int32 SUM over 2^24 elements per GPU (reduction.cpp:665 default).
Step    : ONE full reduction to a scalar.
          N = 1 : one launch of the sm_100a kernel family on one of several rotating inputs (total > 4x L2, so every
                  step streams from HBM, never from the 126 MB L2).
          N > 1 : weak scaling — every GPU owns 2^24 elements; one fused kernel per GPU reduces the local shard and
                  combines the P partials at root over NVLink (b2r_reduce_scalar_sharded; the vendored simpleMultiGPU
                  does this with a host loop, simpleMultiGPU.cpp:264-293). value = N * bytes * K / max-over-ranks time.
e2e     : the same metric through the public Python API on HOST data: every step copies its 64 MiB input from pinned
          host memory (chunked, copy engine overlapped with the reduction kernels) and reads the scalar back.
Reference arm (--impl reference): the UNMODIFIED reference sample built from baseline/_ref (see baseline/README.md),
          run through its own CLI; its stock code path times 100 iterations on ONE L2-resident buffer.

Prints exactly one JSON line on rank 0.
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import time

REPO = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, REPO)



def _workloads():
    """models/workloads.py (the reference's workloads as data) loaded by PATH: importing the package would load libb2r.so, and the
    reference arm must run without any of this repository's native code in its process."""
    import importlib.util

    spec = importlib.util.spec_from_file_location("b2r_workloads", os.path.join(REPO, "cuda_mpi_reductions_b200", "models", "workloads.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules["b2r_workloads"] = mod  # dataclasses looks the module up by name
    spec.loader.exec_module(mod)
    return mod


SDK_REDUCTION = _workloads().SDK_REDUCTION

N_ELEMS = SDK_REDUCTION.elements  # the SDK sample's default workload, 2^24 (reduction.cpp:665)
BASELINE_GBS = SDK_REDUCTION.published_gbs[("int", "SUM")]  # CUDA reduction, int, SUM = 90.8413 (mpi/CUdata.txt:6)
METRIC = "reduction_throughput_int32_sum_2p24"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b2r", choices=["b2r", "reference", "torch_nccl"],
                    help="b2r: this library; reference: the unmodified SDK sample; torch_nccl: what a PyTorch user would write (torch reduction "
                         "kernel + NCCL reduce of the scalar) — the library baseline of BASELINE.json config 4")
    ap.add_argument("--e2e-steps", type=int, default=0, help="steps of the end-to-end arm (default min(steps, 40))")
    ap.add_argument("--dtype", default="int32", choices=["int32", "float32", "float64", "bfloat16"])
    ap.add_argument("--op", default="sum", choices=["sum", "min", "max"])
    ap.add_argument("--n", type=int, default=N_ELEMS)
    ap.add_argument("--no-collectives", action="store_true", help="N > 1: skip the reduce/allreduce-vs-NCCL and config-4 sections")
    return ap.parse_args()


def dist_env():
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", str(rank)))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    return rank, local_rank, world


# --------------------------------------------------------------------------------------------- reference arm
def run_reference(args) -> None:
    rank, local_rank, world = dist_env()
    exe = os.path.join(REPO, "baseline", "_ref", "bin", "reduction")
    out = {"impl": "reference"}
    if not os.path.exists(exe):
        build = os.path.join(REPO, "baseline", "build_reference.sh")
        if os.path.exists(build) and os.path.isdir(os.path.join(REPO, "baseline", "_ref", "src")):
            subprocess.run(["bash", build], capture_output=True)
    if not os.path.exists(exe):
        if rank == 0:
            out["unavailable"] = "baseline/_ref/bin/reduction missing (reference sample not installed; see DESIGN.md)"
            print(json.dumps(out), flush=True)
        return
    import torch

    if world > 1:
        import torch.distributed as dist

        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    typ = {"int32": "int", "float32": "float", "float64": "double"}.get(args.dtype)
    if typ is None:
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "the reference has no bf16 path"}), flush=True)
        return
    # The reference's stock path: fixed 100 timed iterations after one warm-up (reduction.cpp:729-738); --steps/--warmup
    # cannot be passed to it, they are reported as what it actually ran. --n is a power of two by its own rule.
    cmd = [exe, f"--method={args.op.upper()}", f"--type={typ}", f"--n={args.n}", f"--device={local_rank}", "--noprompt"]
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    res = subprocess.run(cmd, capture_output=True, text=True, cwd=os.path.dirname(exe))
    wall = time.perf_counter() - t0
    m = re.search(r"Throughput = ([0-9.]+) GB/s, Time = ([0-9.]+) s", res.stdout)
    passed = "PASSED" in res.stdout
    gbs = float(m.group(1)) if m else 0.0
    esz = {"int32": 4, "float32": 4, "float64": 8}[args.dtype]
    # ---- end to end: the same unmodified kernels driven step by step (baseline/ref_e2e.cu): per step H2D of the input from
    # pinned host memory, the sample's two-pass reduction, D2H of the scalar; --steps / --warmup apply here
    e2e_exe = os.path.join(REPO, "baseline", "_ref", "bin", "ref_e2e")
    e2e_gbs, e2e_ms, e2e_ok, e2e_pageable = 0.0, 0.0, False, 0.0
    if os.path.exists(e2e_exe):
        e2e_steps = max(3, min(args.steps, 40))
        base = [e2e_exe, f"--method={args.op.upper()}", f"--type={typ}", f"--n={args.n}", f"--device={local_rank}",
                f"--steps={e2e_steps}", f"--warmup={max(3, min(args.warmup, 5))}"]
        if world > 1:
            dist.barrier()
        r2 = subprocess.run(base, capture_output=True, text=True)
        m2 = re.search(r"REF_E2E gbs=([0-9.]+) ms_per_step=([0-9.]+) .*result_ok=(\d)", r2.stdout)
        if m2:
            e2e_gbs, e2e_ms, e2e_ok = float(m2.group(1)), float(m2.group(2)), m2.group(3) == "1"
        if world > 1:
            dist.barrier()
        r3 = subprocess.run(base + ["--pageable"], capture_output=True, text=True)
        m3 = re.search(r"REF_E2E gbs=([0-9.]+)", r3.stdout)
        e2e_pageable = float(m3.group(1)) if m3 else 0.0
    if world > 1:
        # the reference has no multi-GPU reduction; N ranks each run the unmodified sample on their own shard and the
        # job-level rate is N x the slowest rank's own printed throughput (the cross-GPU combine, a host loop in
        # simpleMultiGPU.cpp:287-291, is not part of its timed region either). Round 1 rebuilt the rate from the sample's
        # five-decimal `Time =` field, which quantised every multi-GPU figure to +-5 %.
        t = torch.tensor([gbs, e2e_gbs, e2e_pageable], device="cuda", dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        ok = torch.tensor([1 if passed else 0, 1 if e2e_ok else 0], device="cuda")
        dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        tm = torch.tensor([e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tm, op=dist.ReduceOp.MAX)
        gbs, e2e_gbs, e2e_pageable = (world * v for v in t.tolist())
        passed, e2e_ok = bool(ok[0].item()), bool(ok[1].item())
        e2e_ms = tm.item()
    sec = world * args.n * esz / (gbs * 1e9) if gbs > 0 else 0.0
    if rank == 0:
        out.update({
            "metric": METRIC, "value": gbs, "unit": "GB/s", "n_gpus": world, "steps": 100, "warmup": 1,
            "ms_per_step": sec * 1e3, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": gbs / BASELINE_GBS if gbs else None, "dtype": args.dtype, "data": "synthetic (rand()&0xFF, the sample's own generator)",
            "config": {"model": "SDK reduction sample (reference, unmodified) recompiled for sm_100", "elements_per_gpu": args.n,
                       "op": args.op, "parallelism": f"dp{world}" if world > 1 else "single-gpu",
                       "timing": "reference's own host stopwatch around sync+2 launches+sync; one L2-resident buffer (stock path)",
                       "multi_gpu": "N x the slowest rank's printed throughput" if world > 1 else None},
            "self_check": "PASSED" if passed else "FAILED", "wall_s": wall,
        })
        if e2e_gbs > 0:
            out["e2e"] = {"value": e2e_gbs, "unit": "GB/s", "h2d_bytes_per_step": args.n * esz, "d2h_bytes_per_step": esz,
                          "ms_per_step": e2e_ms, "self_check": "PASSED" if e2e_ok else "FAILED",
                          "api": "reduction.h sumreduce/minreduce/maxreduce<T> (unmodified reduction_kernel.cu) driven by baseline/ref_e2e.cu: "
                                 "per step cudaMemcpy H2D of the input, two-pass reduction, cudaMemcpy D2H of the scalar (reduction.cpp:725,323-357,377-381)",
                          "host_memory": "pinned (same rule as the repo's own e2e arm)",
                          "value_pageable_host": e2e_pageable, "note_pageable": "the stock sample mallocs its host buffer (reduction.cpp:696)"}
        if not m:
            out["unavailable"] = "reference binary produced no throughput line: " + (res.stdout[-200:] + res.stderr[-200:]).replace("\n", " ")
        print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- our arm
def run_b2r(args) -> None:
    import torch

    import cuda_mpi_reductions_b200 as b2r
    from cuda_mpi_reductions_b200.utils import ClockSampler, rotating_copies

    rank, local_rank, world = dist_env()
    if args.gpus != world and world > 1 and rank == 0:
        print(f"warning: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE", file=sys.stderr)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (run it through gpurun)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dtype = getattr(torch, args.dtype)
    esz = torch.empty((), dtype=dtype).element_size()
    nbytes = args.n * esz
    steps, warmup = args.steps, max(args.warmup, 3)
    # started now, long before the timed region: the first NVML query of a process takes longer than 20 steps do
    sampler = ClockSampler(local_rank).start() if rank == 0 else None

    comm = None
    do_sections = world > 1 and not args.no_collectives
    if world > 1:
        import torch.distributed as dist

        from cuda_mpi_reductions_b200.parallel import Communicator

        dist.init_process_group("nccl", device_id=dev)
        # symmetric_only: every buffer handed to reduce/allreduce below comes from the symmetric heap, so nothing is staged
        # or chunked; the heap holds the two 1 GiB buffers of the collectives section
        comm = Communicator.from_torch_distributed(heap_bytes=(2 << 30) + (64 << 20) if do_sections else 0, symmetric_only=do_sections,
                                                   timeout_ms=30000)

    copies = rotating_copies(nbytes)
    big = torch.empty(copies * args.n, dtype=dtype, device=dev)
    inputs = [big[c * args.n:(c + 1) * args.n] for c in range(copies)]
    for c, t in enumerate(inputs):
        b2r.fill_device(args.n, dtype, seed=1000 * rank + c, mode="sdk", out=t)
    out = torch.empty((), dtype=dtype, device=dev)
    ws = b2r.Workspace(dev)
    # everything from here on runs on one dedicated (non-default) stream: CUDA graphs cannot be captured on the legacy stream
    torch.cuda.synchronize()
    side = torch.cuda.Stream(dev)
    torch.cuda.set_stream(side)

    # pdl=PDL_INPUT_READY: programmatic dependent launch with early streaming — the inputs were filled long before the loop,
    # so step i+1 may begin streaming its (complete, independent) input while step i is draining; every step still reads its
    # whole input from HBM and finishes its own scalar.
    cfg = b2r.launch_config(pdl=b2r.PDL_INPUT_READY)
    host_path = "regular method"
    if comm is None:
        opc = {"sum": 0, "min": 1, "max": 2}[args.op]

        def step(i):
            b2r.reduce(inputs[i % copies], opc, out=out, workspace=ws, cfg=cfg)

        def step_plain(i):
            b2r.reduce(inputs[i % copies], opc, out=out, workspace=ws)
    else:
        def step_regular(i):
            comm.reduce_scalar_sharded(inputs[i % copies], args.op, root=0, out=out, cfg=cfg)

        # Persistent form of the same call: arguments converted to C types once, one ctypes call per step (~1.5 us less host
        # time per launch; the GPU side of a step is 9 us). Falls back to the regular method if anything about it is off.
        step = step_regular
        if os.environ.get("B2R_BENCH_PREPARED", "1") != "0":
            try:
                step, _ = comm.prepare_reduce_scalar_sharded(inputs, args.op, root=0, out=out, cfg=cfg)
                host_path = "prepared call (Communicator.prepare_reduce_scalar_sharded)"
            except Exception as e:  # noqa: BLE001
                print(f"[rank {rank}] prepared call unavailable ({e}); using the regular method", file=sys.stderr)
                step = step_regular

    # correctness of what is being timed (cheap: one step, on the first of the rotating inputs, against the CPU oracle)
    def verified() -> tuple[bool, str]:
        out.zero_()
        step(0)
        torch.cuda.synchronize()
        if comm is None:
            want = b2r.reduce_oracle(inputs[0].cpu(), args.op)
            return out.cpu().item() == want.item(), f"got {out.item()} want {want.item()}"
        import torch.distributed as dist

        local = b2r.reduce_oracle(inputs[0].cpu(), args.op).to(dev).to(torch.int64 if dtype == torch.int32 else dtype)
        dist.all_reduce(local, op={"sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX}[args.op])
        verdict = torch.ones((), dtype=torch.int32, device=dev)
        msg = ""
        if rank == 0:
            w = local.item()
            if dtype == torch.int32:
                w = (w + 2**31) % 2**32 - 2**31
                good = out.item() == w
            else:
                good = abs(out.item() - w) <= 1e-5 * abs(w) + 1e-6
            msg = f"got {out.item()} want {w}"
            verdict.fill_(1 if good else 0)
        dist.broadcast(verdict, src=0)  # every rank must take the same decision below
        return bool(verdict.item()), msg

    ok, msg = verified()
    if not ok and comm is not None and step is not step_regular:
        print(f"[rank {rank}] prepared call failed the self-check ({msg}); using the regular method", file=sys.stderr)
        step, host_path = step_regular, "regular method (prepared call failed its self-check)"
        ok, msg = verified()
    assert ok, f"self-check failed: {msg}"

    if sampler:
        sampler.wait_for_samples(1, 3.0)
    for i in range(warmup):
        step(i)
    torch.cuda.synchronize()
    # The K timed steps are captured once in a CUDA graph and the timed region is ONE replay of it: at 9 us per step the
    # host launch path (Python + ctypes, 2-4 us per call, with jitter) otherwise decides a 20-step measurement. The graph
    # holds exactly K kernel nodes of ours, with the programmatic (PDL) edges between them; the scalar-sharded reduce numbers
    # its calls on the device, so the multi-GPU flavour replays correctly too. B2R_BENCH_GRAPH=0 times the eager loop instead.
    graph, launch_mode = None, "eager loop (one Python call per step)"
    if os.environ.get("B2R_BENCH_GRAPH", "1") != "0":
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=side):
                for i in range(steps):
                    step(i)
            graph.replay()  # untimed: instantiation / first-replay cost stays out of the timed region
            torch.cuda.synchronize()
            launch_mode = f"one CUDA graph replay of the {steps} steps (captured through the public API)"
        except Exception as e:  # noqa: BLE001
            print(f"[rank {rank}] graph capture unavailable ({e}); timing the eager loop", file=sys.stderr)
            graph = None
    if world > 1:
        flag = torch.tensor([1 if graph is not None else 0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if not flag.item():
            graph, launch_mode = None, "eager loop (one Python call per step)"
        dist.barrier()
    torch.cuda.synchronize()

    def run_timed_steps():
        if graph is not None:
            graph.replay()
        else:
            for i in range(steps):
                step(i)

    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    short_region = steps <= 500
    if sampler and short_region:
        sampler.pause(True)  # no NVML call may sit between ev0 and the launches (see ClockSampler.pause)
        time.sleep(0.003)
    if comm is not None:
        comm.barrier()  # device-side barrier on the launching stream: every GPU passes ev0 within a microsecond or two of its peers
    t_start = time.perf_counter()
    ev0.record()
    run_timed_steps()
    ev1.record()
    torch.cuda.synchronize()
    t_end = time.perf_counter()
    if sampler:
        sampler.pause(False)
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    # Clock record: the timed region is shorter than one NVML poll at --steps 20, so the SAME step loop keeps running,
    # untimed, for >= 100 ms while the sampler collects; the record covers the timed region plus that repeat.
    t_rep = time.perf_counter()
    rep_steps = 0
    while time.perf_counter() - t_rep < 0.12:
        for _ in range(10):
            run_timed_steps()
        rep_steps += 10 * steps
        torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    clocks = sampler.summary(t_start, time.perf_counter()) if sampler else None
    if clocks is not None:
        clocks["window"] = f"timed region + {rep_steps} untimed repeats of the same step ({clocks.get('window')})"
    ms_plain = ms_nopdl = ms_synced = None
    if comm is None:
        from cuda_mpi_reductions_b200.utils import time_steps

        # the same loop as a caller who promises nothing about the inputs gets it: default launch = safe programmatic launch
        # (waits for the previous kernel before its first load) ...
        ms_plain = time_steps(step_plain, steps, warmup)
        # ... with programmatic launch switched off altogether (classic stream order) ...
        cfg_off = b2r.launch_config(pdl=b2r.PDL_OFF)
        ms_nopdl = time_steps(lambda i: b2r.reduce(inputs[i % copies], opc, out=out, workspace=ws, cfg=cfg_off), steps, warmup)
        # ... and timed the reference's way (reduction.cpp:319-374): host stopwatch around sync + launch + sync, per iteration
        torch.cuda.synchronize()
        acc_s = 0.0
        for i in range(steps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            step_plain(i)
            torch.cuda.synchronize()
            acc_s += time.perf_counter() - t0
        ms_synced = acc_s * 1e3

    # ---- end-to-end arm: host (pinned) -> device copy of every step's input + scalar back on the host
    e2e_steps = args.e2e_steps or max(3, min(steps, 40))
    host_copies = 2
    # the pinned inputs of this rank belong on the socket its GPU hangs off (8 ranks sharing the inter-socket link halved the
    # per-GPU rate in round 1); best effort, B2R_NUMA_BIND=0 disables it
    from cuda_mpi_reductions_b200.utils.numa import bind_to_gpu_node

    numa = bind_to_gpu_node(local_rank)
    host_in = [b2r.fill_host(args.n, dtype, seed=77 + 1000 * rank + c, mode="sdk", pin_memory=True) for c in range(host_copies)]
    if comm is None:
        def e2e_step(i):
            return b2r.reduce_from_host(host_in[i % host_copies], args.op, device=dev, workspace=ws)
    else:
        def e2e_step(i):
            return comm.reduce_scalar_sharded_from_host(host_in[i % host_copies], args.op, root=0)
    want_e2e = b2r.reduce_oracle(host_in[0], args.op)
    for i in range(3):
        r = e2e_step(i)
    if comm is None:
        assert e2e_step(0).item() == want_e2e.item(), "e2e self-check failed"
    else:
        # the timed path checked against the CPU oracle of every rank's shard (combined with NCCL, like the headline check)
        got = e2e_step(0)
        local = want_e2e.to(dev).to(torch.int64 if dtype == torch.int32 else dtype)
        dist.all_reduce(local, op={"sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX}[args.op])
        verdict = torch.ones((), dtype=torch.int32, device=dev)
        if rank == 0:
            w = local.item()
            if dtype == torch.int32:
                w = (w + 2**31) % 2**32 - 2**31
                good = got.item() == w
            else:
                good = abs(got.item() - w) <= 1e-5 * abs(w) + 1e-6
            verdict.fill_(1 if good else 0)
        dist.broadcast(verdict, src=0)
        assert bool(verdict.item()), "multi-GPU e2e self-check failed"
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        r = e2e_step(i)  # synchronous: returns the scalar on the host
    torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = t.item()
    chunk_bytes = int(os.environ.get("B2R_H2D_CHUNK_MB", "8")) << 20
    nchunks = -(-nbytes // chunk_bytes)
    acc_sz = 8 if dtype == torch.float64 else 4
    d2h_per_step = nchunks * acc_sz if comm is None else acc_sz  # single GPU: chunk scalars; multi GPU: the one scalar at root

    # ---- N > 1: what BASELINE.json names for several GPUs — reduce / allreduce bus bandwidth against NCCL, and config 4
    sections = {}
    if do_sections:
        from cuda_mpi_reductions_b200.utils.bench_sections import collectives_section, config4_section

        for name, fn in (("collectives", collectives_section), ("config4", config4_section)):
            try:
                sections[name] = fn(comm, dist, dev, rank, world)
            except Exception as e:  # noqa: BLE001  (a failed section must not take the headline line with it)
                sections[name] = {"error": f"{type(e).__name__}: {e}"}
                print(f"[rank {rank}] section {name} failed: {e}", file=sys.stderr)
        dist.barrier()

    if sampler:
        sampler.stop()
    if rank == 0:
        total_bytes = world * nbytes
        value = total_bytes * steps / (ms * 1e-3) * 1e-9
        e2e_value = total_bytes * e2e_steps / e2e_s * 1e-9
        peaks = {}
        try:
            with open(os.path.join(REPO, "MEASURED_PEAKS.json")) as f:
                peaks = json.load(f)
        except Exception:
            pass
        hbm = peaks.get("hbm_gbs", 6650.0)
        line = {
            "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": value / BASELINE_GBS,
            "dtype": args.dtype, "data": "synthetic (counter-based generator, SDK-style values 0..255)",
            "config": {"model": "SDK reduction sample workload (single-pass scalar reduction)", "elements_per_gpu": args.n, "op": args.op,
                       "global_elements": world * args.n, "parallelism": f"dp{world} (sharded scalar reduce to root over NVLink)" if world > 1 else "single-gpu",
                       "cache_control": f"{copies} rotating inputs of {nbytes >> 20} MiB (total {copies * nbytes >> 20} MiB > 4x L2)",
                       "geometry": b2r.plan(args.n, dtype, args.op, device=local_rank, cfg=cfg),
                       "launch": "programmatic dependent launch between consecutive steps (cfg pdl=PDL_INPUT_READY: inputs complete before the loop)", "host_path": host_path, "launch_mode": launch_mode},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": "GB/s", "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": d2h_per_step,
                    "steps": e2e_steps, "ms_per_step": e2e_s * 1e3 / e2e_steps, "self_check": "PASSED",
                    "api": "cuda_mpi_reductions_b200.reduce_from_host(pinned_tensor)" if comm is None else "Communicator.reduce_scalar_sharded_from_host(pinned_tensor)",
                    "pipeline": f"{nchunks} chunks of {chunk_bytes >> 20} MiB: copy engine overlapped with the reduction kernels (native, b2r_reduce{'_scalar_sharded' if comm else ''}_from_host)",
                    "numa_rank0": numa},
            "gpu_launches": steps * world,  # one kernel of ours per GPU per step inside the timed region
            "gpu_launches_per_gpu": steps,
            "hbm_fraction_of_measured": value / world / hbm,
            "impl": "b2r",
        }
        if ms_plain is not None:
            gbs = lambda m: total_bytes * steps / (m * 1e-3) * 1e-9  # noqa: E731
            line["value_plain_stream_order"] = gbs(ms_plain)
            line["value_no_programmatic_launch"] = gbs(ms_nopdl)
            line["value_host_synced"] = gbs(ms_synced)
            line["variants"] = {
                "value": "cfg pdl=PDL_INPUT_READY: the caller declares the inputs complete; consecutive launches overlap their streaming",
                "value_plain_stream_order": "default launch, no promise from the caller: safe programmatic launch (waits for the previous kernel "
                                            "before its first load; L2 prefetch of the slice head while waiting), device events",
                "value_no_programmatic_launch": "cfg pdl=PDL_OFF: classic stream order, device events",
                "value_host_synced": "default launch timed the reference's way: host stopwatch around sync + launch + sync per iteration (reduction.cpp:319-374)",
                "geometry_plain": b2r.plan(args.n, dtype, args.op, device=local_rank)}
        line.update(sections)
        print(json.dumps(line), flush=True)
    if world > 1:
        comm.close()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- library baseline arm
def run_torch_nccl(args) -> None:
    """Same metric, same shapes, same timing rules as run_b2r, done with libraries only: torch's own reduction kernel on
    every GPU, then (N > 1) an NCCL reduce of the per-GPU scalars to rank 0. None of this repository's kernels run here."""
    import torch

    from cuda_mpi_reductions_b200.utils import rotating_copies

    rank, local_rank, world = dist_env()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (run it through gpurun)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    dtype = getattr(torch, args.dtype)
    nbytes = args.n * torch.empty((), dtype=dtype).element_size()
    steps, warmup = args.steps, max(args.warmup, 3)
    if world > 1:
        import torch.distributed as dist

        dist.init_process_group("nccl", device_id=dev)
    copies = rotating_copies(nbytes)
    gen = torch.Generator(device=dev).manual_seed(1000 * rank)
    big = torch.randint(0, 256, (copies * args.n,), dtype=torch.int32, device=dev, generator=gen).to(dtype)
    inputs = [big[c * args.n:(c + 1) * args.n] for c in range(copies)]
    if args.op == "sum":  # accumulate like the library does: wider than the input for int32 / bf16
        acc_dtype = torch.int64 if dtype == torch.int32 else (torch.float32 if dtype == torch.bfloat16 else dtype)
        out = torch.empty((), dtype=acc_dtype, device=dev)

        def local(x):
            torch.sum(x, dim=0, dtype=acc_dtype, out=out)
    else:
        out = torch.empty((), dtype=dtype, device=dev)
        pick = torch.amin if args.op == "min" else torch.amax

        def local(x):
            pick(x, dim=0, out=out)
    red_op = None
    if world > 1:
        red_op = {"sum": dist.ReduceOp.SUM, "min": dist.ReduceOp.MIN, "max": dist.ReduceOp.MAX}[args.op]

    def step(i):
        local(inputs[i % copies])
        if world > 1:
            dist.reduce(out, dst=0, op=red_op)

    for i in range(warmup):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(steps):
        step(i)
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    if world > 1:
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = t.item()
    if rank == 0:
        total_bytes = world * nbytes
        print(json.dumps({
            "impl": "torch_nccl", "metric": METRIC, "value": total_bytes * steps / (ms * 1e-3) * 1e-9, "unit": "GB/s", "n_gpus": world,
            "steps": steps, "warmup": warmup, "ms_per_step": ms / steps, "higher_is_better": True, "scaling": "weak", "dtype": args.dtype,
            "data": "synthetic (torch.randint 0..255)",
            "config": {"model": "torch reduction kernel per GPU + ncclReduce of the scalars to rank 0", "elements_per_gpu": args.n, "op": args.op,
                       "cache_control": f"{copies} rotating inputs of {nbytes >> 20} MiB"},
        }), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.gpus > 1 and "WORLD_SIZE" not in os.environ:
        # convenience: `python bench.py --gpus N` without a launcher re-executes itself under torchrun (the driver starts
        # multi-GPU runs through torch.distributed.run itself, in which case WORLD_SIZE is already set)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", os.environ.get("MASTER_PORT", "29517"), os.path.abspath(__file__)] + sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args)
    elif args.impl == "torch_nccl":
        run_torch_nccl(args)
    else:
        run_b2r(args)


if __name__ == "__main__":
    main()
