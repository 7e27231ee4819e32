"""Keep a GPU process on the CPU cores (and therefore the host memory) of the GPU's own NUMA
node.  Only the host-buffer path cares: pinned staging memory is first-touched by the
calling thread, and a PCIe DMA that crosses the socket interconnect is slower.
"""

from __future__ import annotations

import os
from pathlib import Path


def _parse_cpulist(text: str) -> set[int]:
    cpus: set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_local_cpus(device_index: int) -> set[int]:
    """CPUs attached to the NUMA node of CUDA device ``device_index`` (empty if unknown)."""
    try:
        import torch

        props = torch.cuda.get_device_properties(device_index)
        bdf = f"{props.pci_domain_id:04x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        path = Path("/sys/bus/pci/devices") / bdf / "local_cpulist"
        if path.exists():
            return _parse_cpulist(path.read_text())
    except Exception:
        pass
    return set()


def bind_to_gpu_numa(device_index: int) -> dict:
    """Restrict this process to the allowed CPUs local to the GPU.  Returns what was done."""
    allowed = os.sched_getaffinity(0)
    local = gpu_local_cpus(device_index)
    target = allowed & local
    info = {"allowed": len(allowed), "gpu_local": len(local), "bound": 0}
    if target and target != allowed:
        try:
            os.sched_setaffinity(0, target)
            info["bound"] = len(target)
        except OSError:
            pass
    return info
