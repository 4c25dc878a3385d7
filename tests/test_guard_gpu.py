"""GPU: out-of-bounds detection with guard pages (b2r_guard_alloc) — the library's stand-in for a memory checker
(reference: cutil's emulation-mode bank checker, bank_checker.h:25-60; compute-sanitizer is closed on the GPU pool).

Every kernel variant of the single-GPU family reduces buffers whose LAST byte is the last mapped byte of an allocation (and
buffers whose first byte is the first mapped one): one vector read too far, one mis-sized bulk copy and the launch dies with an
illegal-address fault. The work runs in a child process, because such a fault poisons the CUDA context for the rest of the
process; the parent only looks at the child's verdict."""
import os
import subprocess
import sys
import textwrap

import pytest

pytestmark = pytest.mark.gpu
REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

CHILD = textwrap.dedent("""
    import ctypes as C, sys
    import torch
    sys.path.insert(0, %r)
    import cuda_mpi_reductions_b200 as b2r
    from cuda_mpi_reductions_b200 import _lib

    RAW = {torch.int32: "<i4", torch.float32: "<f4", torch.float64: "<f8", torch.bfloat16: "<u2"}

    class Buf:
        def __init__(self, ptr, numel, typestr):
            self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2}

    def guarded(host, at_end):
        nbytes = host.numel() * host.element_size()
        p = C.c_void_p()
        _lib.check(_lib.LIB.b2r_guard_alloc(nbytes, 0, int(at_end), C.byref(p)), "b2r_guard_alloc")
        t = torch.as_tensor(Buf(p.value, host.numel(), RAW[host.dtype]), device="cuda")
        if host.dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        t.copy_(host)
        return t, p

    torch.cuda.set_device(0)
    ws = b2r.Workspace()
    checked = 0
    cfgs = [None, {"variant": "ldg256"}, {"variant": "tma"}, {"variant": "tma", "pdl": 2}, {"pdl": -1}, {"dyn_pct": 60}, {"cluster": 2},
            {"variant": "ldg128", "unroll": 8}]
    for dtype in (torch.int32, torch.float64, torch.bfloat16):
        esz = torch.empty((), dtype=dtype).element_size()
        for nbytes in (16, 4096 + esz, 65536 - esz, (1 << 20) + 3 * esz, (9 << 20) + 48, (40 << 20) - esz):
            n = nbytes // esz
            host = b2r.fill_host(n, dtype, seed=nbytes, mode="sdk")
            want = b2r.reduce_oracle(host, "max").item()
            for at_end in (1, 0):
                x, p = guarded(host, at_end)
                for cfg in cfgs:
                    for rep in range(2):  # twice: the second launch runs behind the first one's programmatic-launch prefetch
                        got = b2r.reduce(x, "max", workspace=ws, cfg=cfg)
                    assert got.item() == want, (dtype, nbytes, at_end, cfg, got.item(), want)
                    checked += 1
                torch.cuda.synchronize()  # a fault surfaces here at the latest
                del x
                _lib.check(_lib.LIB.b2r_guard_free(p), "b2r_guard_free")
    print("GUARD_OK", checked)
""") % REPO


def test_single_gpu_family_never_touches_a_byte_outside_its_input():
    res = subprocess.run([sys.executable, "-c", CHILD], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0 and "GUARD_OK" in res.stdout, (res.stdout[-800:] + res.stderr[-1500:])
