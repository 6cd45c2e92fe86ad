"""NUMA placement of the host-side threads of one rank (hard part 1).

The end-to-end limit of the ingest path is host memory traffic: every payload byte is
written into pinned memory by the pack threads and read back by the GPU's DMA engine.  On a
multi-socket host both should stay on the memory controller the GPU hangs off: the threads
that fill a rank's staging slots (the producer threads and the native copy pool) are pinned to
the cores of the GPU's NUMA node, and ranks whose GPUs share a node split its cores.  Pinned
slot memory is allocated -- and first touched -- by those threads' process, so it lands on the
same node under the default first-touch policy.

Everything here reads sysfs and degrades to "no pinning" when the topology is not visible
(containers without /sys/bus/pci, single-node hosts, non-Linux).
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional


def _parse_cpulist(text: str) -> List[int]:
    cpus: List[int] = []
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.extend(range(int(a), int(b) + 1))
        else:
            cpus.append(int(part))
    return cpus


def numa_nodes() -> Dict[int, List[int]]:
    """{node: cpus} from /sys/devices/system/node; {} when not visible."""
    base = "/sys/devices/system/node"
    out: Dict[int, List[int]] = {}
    try:
        for name in os.listdir(base):
            if name.startswith("node") and name[4:].isdigit():
                with open(os.path.join(base, name, "cpulist")) as f:
                    out[int(name[4:])] = _parse_cpulist(f.read())
    except OSError:
        return {}
    return out


def gpu_numa_node(device_index: int) -> Optional[int]:
    """NUMA node of a CUDA device (its PCI function's numa_node), None when unknown."""
    try:
        import torch
        prop = torch.cuda.get_device_properties(device_index)
        bdf = "%04x:%02x:%02x.0" % (prop.pci_domain_id, prop.pci_bus_id, prop.pci_device_id)
        with open(f"/sys/bus/pci/devices/{bdf}/numa_node") as f:
            node = int(f.read().strip())
        return node if node >= 0 else None
    except Exception:
        return None


def rank_cpus(device_index: int, local_rank: Optional[int] = None,
              local_world: Optional[int] = None) -> Optional[List[int]]:
    """Cores for the host threads of the rank that drives ``device_index``: the cores of the GPU's
    NUMA node that this process may use, divided among the local ranks whose GPUs sit on the same
    node (rank i is assumed to drive GPU i, the torchrun convention).  None = leave threads alone."""
    try:
        allowed = sorted(os.sched_getaffinity(0))
    except AttributeError:  # pragma: no cover
        return None
    if local_rank is None:
        local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if local_world is None:
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", "1"))
    nodes = numa_nodes()
    node = gpu_numa_node(device_index)
    if node is None or node not in nodes or len(nodes) < 2:
        # one memory domain (or unknown): only divide the cores among the local ranks
        if local_world <= 1:
            return None
        share = [c for k, c in enumerate(allowed) if k * local_world // len(allowed) == local_rank]
        return share or None
    cpus = [c for c in nodes[node] if c in set(allowed)]
    if not cpus:
        return None
    # local ranks on the same node, in rank order
    peers = [r for r in range(local_world)
             if (gpu_numa_node(r) if r != local_rank else node) == node] if local_world > 1 else [local_rank]
    if local_rank not in peers:
        peers.append(local_rank)
        peers.sort()
    k, m = peers.index(local_rank), len(peers)
    share = cpus[k * len(cpus) // m:(k + 1) * len(cpus) // m]
    return share or cpus


def pin_current_thread(cpus: Optional[List[int]]) -> bool:
    """Pin the calling thread (not the process) to ``cpus``."""
    if not cpus:
        return False
    try:
        import threading
        os.sched_setaffinity(threading.get_native_id(), cpus)
        return True
    except (AttributeError, OSError):
        return False
