"""NUMA placement of the pinned host buffers of the host-fed field query (field.HostFieldQuery, field._field_query_host).

The host-fed path moves 28 bytes per point over PCIe (12 up, 16 down) and is bound by the host side at N > 1: eight ranks
share the host's memory controllers and, when a rank's pinned pages sit on the other socket, its inter-socket link as
well.  Linux places pages on the node of the thread that first touches them (cudaHostAlloc touches them while locking), so
binding the rank's thread to the CPUs of its GPU's NUMA node BEFORE the first pinned allocation makes the buffers (and the
thread that launches the copies) local to the GPU's PCIe root.  The reference has no counterpart: it feeds its queries
from device tensors (nav/quad_plot.py:224-250) and runs one process on one GPU.

    with hostmem.numa_local(device):          # affinity restored on exit; pages allocated inside stay where they are
        x_host = torch.empty(B, 3).pin_memory()
        hq = HostFieldQuery(model, B)

Everything here is host-side bookkeeping read from sysfs; where sysfs has no answer (numa_node = -1 in a VM, no
/sys/bus/pci) the context manager does nothing and `describe()` says so.
"""
import contextlib
import os

import torch


def _parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        if "-" in part:
            a, b = part.split("-")
            cpus.update(range(int(a), int(b) + 1))
        else:
            cpus.add(int(part))
    return cpus


def pci_address(device):
    """'dddd:bb:dd.0' of a CUDA device (torch's view of it, so CUDA_VISIBLE_DEVICES is honoured), or None."""
    try:
        p = torch.cuda.get_device_properties(device)
        return f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
    except Exception:
        return None


def gpu_numa_node(device, sysfs="/sys"):
    """NUMA node of the GPU's PCIe root according to sysfs, or None when unknown (-1, no sysfs entry, no GPU)."""
    addr = pci_address(device)
    if addr is None:
        return None
    try:
        with open(os.path.join(sysfs, "bus/pci/devices", addr, "numa_node")) as f:
            node = int(f.read().strip())
    except (OSError, ValueError):
        return None
    return node if node >= 0 else None


def node_cpus(node, sysfs="/sys"):
    """CPUs of a NUMA node that this process may run on (the node's cpulist intersected with the current affinity)."""
    try:
        with open(os.path.join(sysfs, f"devices/system/node/node{node}/cpulist")) as f:
            cpus = _parse_cpulist(f.read())
    except (OSError, ValueError):
        return set()
    return cpus & os.sched_getaffinity(0)


def describe(device, sysfs="/sys"):
    """What numa_local(device) would do, for the bench line: {'pci', 'node', 'cpus', 'bound'}."""
    node = gpu_numa_node(device, sysfs)
    cpus = node_cpus(node, sysfs) if node is not None else set()
    return {"pci": pci_address(device), "node": node, "cpus": len(cpus), "bound": bool(cpus)}


@contextlib.contextmanager
def numa_local(device, sysfs="/sys"):
    """Run the body with the calling thread bound to the CPUs of `device`'s NUMA node (no-op when sysfs does not say).
    Threads started inside inherit the binding; the previous affinity of the calling thread is restored on exit."""
    node = gpu_numa_node(device, sysfs)
    cpus = node_cpus(node, sysfs) if node is not None else set()
    if not cpus:
        yield None
        return
    before = os.sched_getaffinity(0)
    os.sched_setaffinity(0, cpus)
    try:
        yield node
    finally:
        os.sched_setaffinity(0, before)
