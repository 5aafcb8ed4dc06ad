"""CPU / memory affinity of a rank: run on the NUMA node its GPU hangs off, so that the pinned host buffers the input
pipeline allocates (first touch) and the threads that fill them sit next to that GPU's PCIe root port.

The reference leaves placement to Ray's worker scheduling (``ScalingConfig``, /root/reference/my_ray_module.py:262-267);
here a rank is one process per GPU and binds itself.
"""
from __future__ import annotations

import os
from typing import Dict, Optional


def _parse_cpulist(text: str):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def gpu_numa_node(device_index: int) -> Optional[int]:
    """NUMA node of a CUDA device from sysfs (None when the platform does not say, e.g. inside some VMs)."""
    try:
        import torch
        bus = torch.cuda.get_device_properties(device_index)
        pci = f"{bus.pci_domain_id:04x}:{bus.pci_bus_id:02x}:{bus.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{pci}/numa_node") as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def bind_to_gpu_numa(device_index: int) -> Dict[str, object]:
    """Restrict this process to the CPUs of its GPU's NUMA node (memory follows by first touch). Returns what was done."""
    info: Dict[str, object] = {"numa_node": None, "cpus": len(os.sched_getaffinity(0)), "bound": False}
    node = gpu_numa_node(device_index)
    info["numa_node"] = node
    if node is None:
        return info
    try:
        with open(f"/sys/devices/system/node/node{node}/cpulist") as f:
            cpus = _parse_cpulist(f.read()) & os.sched_getaffinity(0)
        if cpus:
            os.sched_setaffinity(0, cpus)
            info.update(cpus=len(cpus), bound=True)
    except Exception:
        pass
    return info


def bind_rank_cores(local_rank: int, local_world: int) -> Dict[str, object]:
    """Give this rank its own contiguous slice of the CPUs the process may run on (what ``rtdc-launch --bind-cores`` does for
    workers it starts): under an external launcher (torchrun) eight unpinned ranks with their loader / checkpoint / report
    threads migrate across all cores, and a training thread that loses its core for a moment stalls every GPU at the next
    gradient barrier. No-op when there are fewer CPUs than ranks or ``RTDC_NO_BIND`` is set."""
    info: Dict[str, object] = {"cpus": len(os.sched_getaffinity(0)), "bound": False}
    if os.environ.get("RTDC_NO_BIND") or local_world <= 1:
        return info
    try:
        avail = sorted(os.sched_getaffinity(0))
        per = len(avail) // local_world
        if per < 1:
            return info
        mine = avail[local_rank * per:(local_rank + 1) * per]
        os.sched_setaffinity(0, mine)
        info.update(cpus=len(mine), bound=True, first=mine[0])
    except Exception:
        pass
    return info
