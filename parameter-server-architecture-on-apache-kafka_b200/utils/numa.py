"""Keep a rank's threads and its pinned staging memory on the NUMA node its GPU hangs off.

One process per GPU streams its batches out of pinned host memory every clock (28 GB/s per GPU in the end-to-end
benchmark).  On a two-socket box a rank whose staging memory was first touched on the other socket makes every one
of those reads cross the socket interconnect, which eight ranks together saturate; a rank scheduled on the far socket
also pays for it on every launch.  Before it allocates anything pinned, a rank therefore restricts itself to the CPUs
of its GPU's node (``sched_setaffinity``; threads started later inherit it) and prefers that node's memory
(``set_mempolicy(MPOL_PREFERRED)``, which also works when the container's CPUs all sit on the other socket).

Everything here is best effort: no sysfs entry, a single node, a CPU set the container does not allow, a kernel that
refuses the policy — each of them leaves the process as it was.  ``PSB_NUMA_BIND=0`` switches it off.
"""
from __future__ import annotations

import ctypes
import os
import re
from typing import Dict, Optional, Set

MPOL_DEFAULT, MPOL_PREFERRED = 0, 1
SYS_SET_MEMPOLICY = {"x86_64": 238, "aarch64": 237}


def parse_cpulist(text: str) -> Set[int]:
    """'0-3,8,10-11' -> {0, 1, 2, 3, 8, 10, 11} (the format of sysfs cpulist files)."""
    cpus: Set[int] = set()
    for part in text.strip().split(","):
        if not part:
            continue
        lo, _, hi = part.partition("-")
        cpus.update(range(int(lo), int(hi or lo) + 1))
    return cpus


def pci_address(domain: int, bus: int, device: int) -> str:
    return f"{domain:04x}:{bus:02x}:{device:02x}.0"


def node_of_pci_device(bdf: str, sysfs: str = "/sys") -> Optional[int]:
    try:
        with open(os.path.join(sysfs, "bus", "pci", "devices", bdf, "numa_node")) as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def online_nodes(sysfs: str = "/sys") -> Set[int]:
    try:
        names = os.listdir(os.path.join(sysfs, "devices", "system", "node"))
    except OSError:
        return set()
    return {int(m.group(1)) for m in (re.fullmatch(r"node(\d+)", n) for n in names) if m}


def cpus_of_node(node: int, sysfs: str = "/sys") -> Set[int]:
    try:
        with open(os.path.join(sysfs, "devices", "system", "node", f"node{node}", "cpulist")) as f:
            return parse_cpulist(f.read())
    except (OSError, ValueError):
        return set()


def prefer_memory_of(node: int) -> bool:
    """set_mempolicy(MPOL_PREFERRED, {node}) for the calling thread (inherited by threads it starts)."""
    nr = SYS_SET_MEMPOLICY.get(os.uname().machine)
    if nr is None or node < 0 or node >= 1024:
        return False
    words = node // 64 + 1
    mask = (ctypes.c_ulong * words)()
    mask[node // 64] = 1 << (node % 64)
    libc = ctypes.CDLL(None, use_errno=True)
    return libc.syscall(nr, MPOL_PREFERRED, mask, words * 64 + 1) == 0


def bind_to_node(node: Optional[int], sysfs: str = "/sys", apply: bool = True) -> Optional[Dict]:
    """Restrict the calling thread to `node`'s CPUs (those the process may use) and prefer its memory.  Returns what
    was done, or None when there is nothing to do (unknown node, or a machine with a single node)."""
    if node is None or len(online_nodes(sysfs)) < 2:
        return None
    allowed = os.sched_getaffinity(0)
    cpus = cpus_of_node(node, sysfs) & allowed
    done = {"node": node, "cpus": len(cpus) if cpus else len(allowed), "cpus_bound": False, "memory_preferred": False}
    if not apply:
        return done
    if cpus and cpus != allowed:
        try:
            os.sched_setaffinity(0, cpus)
            done["cpus_bound"] = True
        except OSError:
            done["cpus"] = len(allowed)
    try:
        done["memory_preferred"] = prefer_memory_of(node)
    except Exception:
        pass
    return done


def bind_to_device_node(device_index: int, sysfs: str = "/sys") -> Optional[Dict]:
    """bind_to_node() for the node of CUDA device `device_index`.  Never raises."""
    if os.environ.get("PSB_NUMA_BIND", "1") == "0":
        return None
    try:
        import torch
        p = torch.cuda.get_device_properties(device_index)
        return bind_to_node(node_of_pci_device(pci_address(p.pci_domain_id, p.pci_bus_id, p.pci_device_id), sysfs), sysfs)
    except Exception:
        return None


__all__ = ["parse_cpulist", "pci_address", "node_of_pci_device", "online_nodes", "cpus_of_node", "prefer_memory_of",
           "bind_to_node", "bind_to_device_node"]
