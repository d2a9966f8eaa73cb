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

Everything here is host-side bookkeeping read from sysfs, with NVML's ideal CPU affinity of the GPU as the second source
(containers often hide the PCI device's numa_node); where neither has an answer (numa_node = -1 in a VM, an NVML set
that covers every CPU) the context manager does nothing and `describe()` says so.
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


def nvml_cpus(device):
    """The driver's own answer (NVML's ideal CPU affinity of the GPU, what `nvidia-smi topo -m` prints), for boxes whose
    sysfs hides the PCI device's numa_node (containers): the CPUs in it that this process may run on, or an empty set.
    An answer that covers the whole current affinity carries no placement information and counts as none."""
    addr = pci_address(device)
    if addr is None:
        return set()
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            h = pynvml.nvmlDeviceGetHandleByPciBusId(addr.encode())
            words = pynvml.nvmlDeviceGetCpuAffinity(h, (max(os.cpu_count() or 64, 64) + 63) // 64)
        finally:
            pynvml.nvmlShutdown()
    except Exception:
        return set()
    cpus = {64 * i + b for i, w in enumerate(words) for b in range(64) if (int(w) >> b) & 1}
    mine = os.sched_getaffinity(0)
    cpus &= mine
    return set() if cpus == mine else cpus


def _local_cpus(device, sysfs):
    node = gpu_numa_node(device, sysfs)
    cpus = node_cpus(node, sysfs) if node is not None else set()
    source = "sysfs" if cpus else None
    if not cpus:
        cpus = nvml_cpus(device)
        source = "nvml" if cpus else None
    return node, cpus, source


def describe(device, sysfs="/sys"):
    """What numa_local(device) would do, for the bench line: {'pci', 'node', 'cpus', 'bound', 'source'}."""
    node, cpus, source = _local_cpus(device, sysfs)
    return {"pci": pci_address(device), "node": node, "cpus": len(cpus), "bound": bool(cpus), "source": source}


@contextlib.contextmanager
def numa_local(device, sysfs="/sys"):
    """Run the body with the calling thread bound to the CPUs of `device`'s NUMA node (no-op when sysfs does not say).
    Threads started inside inherit the binding; the previous affinity of the calling thread is restored on exit."""
    node, cpus, source = _local_cpus(device, sysfs)
    if not cpus:
        yield None
        return
    before = os.sched_getaffinity(0)
    os.sched_setaffinity(0, cpus)
    try:
        yield node if node is not None else -1      # -1: bound from NVML's CPU set, node number unknown
    finally:
        os.sched_setaffinity(0, before)
