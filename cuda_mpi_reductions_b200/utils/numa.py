"""Keep a rank's host threads (and therefore the pages they first touch) on the NUMA node its GPU hangs off.

Why: the end-to-end arm streams every step's input from pinned host memory. One rank alone reaches ~52 GB/s over PCIe;
with eight ranks started by torchrun the figure fell to ~24 GB/s per GPU in round 1, because nothing tied a rank's pinned
buffers to the socket its GPU is attached to — half of the copies crossed the inter-socket link. Linux places a page on
the node of the CPU that first touches it (the driver pins it where it already is), so binding the process to the CPUs of
the GPU's node before allocating is enough; no libnuma needed.

Everything here is best effort: a box without sysfs NUMA information, a container whose cpuset excludes the node, or a
single-node machine simply leaves the affinity alone. Set B2R_NUMA_BIND=0 to switch it off.
"""
from __future__ import annotations

import os


def parse_cpulist(text: str) -> set[int]:
    """'0-3,8,10-11' -> {0,1,2,3,8,10,11} (the format of /sys/devices/system/node/nodeN/cpulist)."""
    cpus: set[int] = set()
    for part in text.strip().split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            lo, hi = part.split("-", 1)
            cpus.update(range(int(lo), int(hi) + 1))
        else:
            cpus.add(int(part))
    return cpus


def pci_address(device_index: int) -> str | None:
    """'0000:1b:00.0' for a CUDA device ordinal (honours CUDA_VISIBLE_DEVICES because it asks the CUDA runtime)."""
    try:
        import torch

        p = torch.cuda.get_device_properties(device_index)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        return None


def gpu_numa_node(pci_addr: str, sysfs_root: str = "/sys") -> int | None:
    try:
        with open(os.path.join(sysfs_root, "bus/pci/devices", pci_addr, "numa_node")) as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None  # -1: the platform does not say


def nvml_cpu_affinity(device_index: int) -> set[int]:
    """The CPUs NVML calls 'ideal' for the device (what `nvidia-smi topo -m` prints as CPU Affinity). Used when sysfs has
    no numa_node for the PCI function, which is the case in the containers this library is benchmarked in."""
    try:
        import pynvml

        pynvml.nvmlInit()
        import torch

        uuid = str(torch.cuda.get_device_properties(device_index).uuid)
        try:
            h = pynvml.nvmlDeviceGetHandleByUUID(("GPU-" + uuid).encode())
        except Exception:
            h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = set()
        for w, mask in enumerate(words):
            for b in range(64):
                if int(mask) >> b & 1:
                    cpus.add(64 * w + b)
        return cpus
    except Exception:
        return set()


def node_cpus(node: int, sysfs_root: str = "/sys") -> set[int]:
    try:
        with open(os.path.join(sysfs_root, f"devices/system/node/node{node}/cpulist")) as f:
            return parse_cpulist(f.read())
    except (OSError, ValueError):
        return set()


def bind_to_gpu_node(device_index: int, sysfs_root: str = "/sys", pci_addr: str | None = None, apply: bool = True) -> dict:
    """Restricts this process to the CPUs of the NUMA node of `device_index`. Returns what was found and done:
    {"pci": ..., "node": ..., "cpus": n, "bound": bool, "why": text}. Never raises."""
    info = {"pci": None, "node": None, "cpus": 0, "bound": False, "why": ""}
    if os.environ.get("B2R_NUMA_BIND", "1") in ("0", "false", "no"):
        info["why"] = "disabled by B2R_NUMA_BIND"
        return info
    addr = pci_addr or pci_address(device_index)
    info["pci"] = addr
    if not addr:
        info["why"] = "no PCI address for the device"
        return info
    node = gpu_numa_node(addr, sysfs_root)
    info["node"] = node
    try:
        allowed = os.sched_getaffinity(0)
    except (AttributeError, OSError):
        info["why"] = "sched_getaffinity unavailable"
        return info
    if node is None:
        # sysfs is silent (containers): ask NVML for the device's CPU affinity, the column `nvidia-smi topo -m` prints
        cpus = nvml_cpu_affinity(device_index) & allowed if sysfs_root == "/sys" else set()
        info["source"] = "nvml cpu affinity"
        if not cpus:
            info["why"] = "neither sysfs nor NVML reports a NUMA placement for the device"
            return info
    else:
        cpus = node_cpus(node, sysfs_root) & allowed
    info["cpus"] = len(cpus)
    if not cpus:
        info["why"] = "none of the node's CPUs are in this process's cpuset"
        return info
    if cpus == allowed:
        info["why"] = "already confined to that node"
        info["bound"] = True
        return info
    if apply:
        try:
            os.sched_setaffinity(0, cpus)
        except OSError as e:
            info["why"] = f"sched_setaffinity failed: {e}"
            return info
    info["bound"] = True
    info["why"] = "bound" if apply else "dry run"
    return info
