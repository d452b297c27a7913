"""Bind a rank to the CPUs next to its GPU (opt-in: ``CCMPI_NUMA_BIND=1``).

With one process per GPU, host->device copies of a training loop read pinned host memory that the rank
itself allocated; if the rank runs on the other socket, every batch crosses the inter-socket link before
it reaches the GPU's PCIe root port. ``mpirun`` would be told ``--bind-to`` / ``--map-by`` for this; the
reference never needed it (README.md:52 runs CPU ranks). Here the rank pins itself: GPU -> PCI address ->
``/sys/bus/pci/devices/<addr>/numa_node`` -> that node's CPU list -> ``sched_setaffinity``. Memory the
process touches afterwards (pinned batches included) is then allocated on that node by first touch.

Everything is best effort: no sysfs entry, node -1, or a cpuset that forbids the CPUs -> nothing changes.
Not measured: the pool's 8-GPU boxes are virtual machines that expose ONE NUMA node (CPU affinity 0-31 and NUMA
node 0 for all eight GPUs, profiles/r1_topo_8gpu.txt), so binding is a no-op there. Hence opt-in.
"""
from __future__ import annotations

import os
from typing import List, Optional


def parse_cpulist(text: str) -> List[int]:
    """``"0-3,8,10-11"`` -> ``[0, 1, 2, 3, 8, 10, 11]``."""
    cpus: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.extend(range(int(lo), int(hi or lo) + 1))
    return cpus


def pci_address(domain: int, bus: int, device: int) -> str:
    return f"{domain:04x}:{bus:02x}:{device:02x}.0"


def node_of_pci_device(addr: str, sysfs: str = "/sys") -> Optional[int]:
    try:
        with open(os.path.join(sysfs, "bus", "pci", "devices", addr, "numa_node")) as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def cpus_of_node(node: int, sysfs: str = "/sys") -> List[int]:
    try:
        with open(os.path.join(sysfs, "devices", "system", "node", f"node{node}", "cpulist")) as f:
            return parse_cpulist(f.read())
    except (OSError, ValueError):
        return []


def bind_to_gpu_node(device_index: int, sysfs: str = "/sys") -> Optional[int]:
    """Restrict this process to the CPUs of the NUMA node ``device_index`` hangs off. Returns the node, or
    ``None`` when nothing was changed."""
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        addr = pci_address(int(p.pci_domain_id), int(p.pci_bus_id), int(p.pci_device_id))
    except Exception:  # noqa: BLE001  no GPU, or a torch build without the PCI fields
        return None
    node = node_of_pci_device(addr, sysfs)
    if node is None:
        return None
    allowed = set(os.sched_getaffinity(0))
    cpus = [c for c in cpus_of_node(node, sysfs) if c in allowed]
    if not cpus:
        return None
    try:
        os.sched_setaffinity(0, cpus)
    except OSError:
        return None
    return node
