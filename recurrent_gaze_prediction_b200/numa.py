"""Bind a rank to the NUMA node of its GPU (best effort; plumbing only).

The end-to-end step is a host->device copy of the fed features at PCIe line rate.  With one process
per GPU and every rank's pinned buffers on NUMA node 0, eight ranks pull ~0.45 TB/s out of one socket's DRAM and across the
inter-socket link; allocating each rank's pinned staging memory on the node its GPU hangs off removes that.  Call
``bind_to_gpu_node(local_rank)`` BEFORE the pinned buffers are allocated (first touch decides the page placement).

Nothing here is required for correctness: on a box with one node, a restricted cpuset, or no sysfs, the function
reports what it found and changes nothing.
"""
from __future__ import annotations

import ctypes
import os


def _read(path):
    try:
        with open(path) as fp:
            return fp.read().strip()
    except OSError:
        return None


def _parse_cpulist(s):
    cpus = set()
    for part in (s or "").split(","):
        part = part.strip()
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def gpu_pci_bus_id(index):
    """'0000:1b:00.0' of CUDA device `index` of this process (honours CUDA_VISIBLE_DEVICES), or None."""
    try:
        import torch
        p = torch.cuda.get_device_properties(index)
        dom, bus, dev = getattr(p, "pci_domain_id", None), getattr(p, "pci_bus_id", None), getattr(p, "pci_device_id", None)
        if bus is not None and dev is not None:
            return "%04x:%02x:%02x.0" % (dom or 0, bus, dev)
    except Exception:
        pass
    try:
        import pynvml
        pynvml.nvmlInit()
        vis = os.environ.get("CUDA_VISIBLE_DEVICES")
        idx = index
        if vis and all(x.strip().isdigit() for x in vis.split(",")):
            idx = int(vis.split(",")[index])
        bid = pynvml.nvmlDeviceGetPciInfo(pynvml.nvmlDeviceGetHandleByIndex(idx)).busId
        bid = bid.decode() if isinstance(bid, bytes) else bid
        return bid.lower()[-12:]
    except Exception:
        return None


def bind_to_gpu_node(local_rank=0):
    """Returns {'node', 'cpus_bound', 'mem_policy', 'note'}."""
    out = {"node": None, "cpus_bound": 0, "mem_policy": "unchanged", "note": ""}
    bid = gpu_pci_bus_id(local_rank)
    if bid is None:
        out["note"] = "no PCI bus id for device %d" % local_rank
        return out
    node = _read("/sys/bus/pci/devices/%s/numa_node" % bid)
    if node is None or int(node) < 0:
        out["note"] = "sysfs reports no NUMA node for %s" % bid
        # Topology unknown (virtualised PCI): at least do not let every rank's pinned staging pages land on the node the
        # ranks happened to start on - interleave new pages over all nodes this process may allocate from.
        nodes = _online_nodes()
        if len(nodes) > 1:
            rc = _set_mempolicy(3, nodes)                      # MPOL_INTERLEAVE
            out["mem_policy"] = ("interleave over nodes %s" % nodes) if rc == 0 else "set_mempolicy(interleave) errno %d" % rc
        return out
    node = int(node)
    out["node"] = node
    cpus = _parse_cpulist(_read("/sys/devices/system/node/node%d/cpulist" % node))
    try:
        allowed = os.sched_getaffinity(0)
        mine = sorted(cpus & set(allowed))
        if mine:
            os.sched_setaffinity(0, mine)
            out["cpus_bound"] = len(mine)
        else:
            out["note"] = "cpuset %s has no CPU of node %d; affinity unchanged" % (_fmt(allowed), node)
    except (AttributeError, OSError) as e:
        out["note"] = "sched_setaffinity: %s" % e
    # prefer this node for new pages (pinned staging buffers): set_mempolicy(MPOL_PREFERRED, 1 << node)
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        nbits = 1024
        mask = (ctypes.c_ulong * (nbits // (8 * ctypes.sizeof(ctypes.c_ulong))))()
        mask[node // (8 * ctypes.sizeof(ctypes.c_ulong))] |= 1 << (node % (8 * ctypes.sizeof(ctypes.c_ulong)))
        SYS_set_mempolicy, MPOL_PREFERRED = 238, 1          # x86_64
        if os.uname().machine == "aarch64":
            SYS_set_mempolicy = 237
        rc = libc.syscall(SYS_set_mempolicy, MPOL_PREFERRED, mask, nbits)
        out["mem_policy"] = "preferred node %d" % node if rc == 0 else "set_mempolicy errno %d" % ctypes.get_errno()
    except Exception as e:       # no libc / syscall filtered
        out["mem_policy"] = "unavailable (%s)" % e
    return out


def _online_nodes():
    s = _read("/sys/devices/system/node/online")
    return sorted(_parse_cpulist(s)) if s else []


def _set_mempolicy(mode, nodes):
    """set_mempolicy(mode, nodemask of `nodes`); returns 0 or errno (never raises)"""
    try:
        libc = ctypes.CDLL(None, use_errno=True)
        nbits = 1024
        wbits = 8 * ctypes.sizeof(ctypes.c_ulong)
        mask = (ctypes.c_ulong * (nbits // wbits))()
        for n in nodes:
            mask[n // wbits] |= 1 << (n % wbits)
        nr = 237 if os.uname().machine == "aarch64" else 238
        rc = libc.syscall(nr, mode, mask, nbits)
        return 0 if rc == 0 else (ctypes.get_errno() or -1)
    except Exception:
        return -1


def _fmt(cpus):
    cpus = sorted(cpus)
    return "%d-%d (%d)" % (cpus[0], cpus[-1], len(cpus)) if cpus else "empty"
