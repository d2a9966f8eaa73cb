"""hostmem.numa_local: NUMA placement of the host-fed query's pinned buffers (no reference counterpart; host-side
bookkeeping over sysfs).  A fake sysfs tree stands in for the box: node lookup, cpulist parsing, the affinity set inside
the context and restored after it, and the no-op cases (numa_node = -1, no entry, CPUs outside the current affinity)."""
import os

import pytest

from nerf_navigation_b200 import hostmem


def _fake_sysfs(tmp_path, addr, node, cpulists):
    d = tmp_path / "bus" / "pci" / "devices" / addr
    d.mkdir(parents=True)
    (d / "numa_node").write_text(f"{node}\n")
    for n, text in cpulists.items():
        nd = tmp_path / "devices" / "system" / "node" / f"node{n}"
        nd.mkdir(parents=True)
        (nd / "cpulist").write_text(text + "\n")
    return str(tmp_path)


def test_cpulist_parsing():
    assert hostmem._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert hostmem._parse_cpulist("\n") == set()


def test_binding_and_restore(tmp_path, monkeypatch):
    mine = sorted(os.sched_getaffinity(0))
    if len(mine) < 2:
        pytest.skip("needs two CPUs in the affinity mask")
    half = mine[: len(mine) // 2]
    addr = "0000:1b:00.0"
    sysfs = _fake_sysfs(tmp_path, addr, 1, {0: "100000-100003", 1: ",".join(map(str, half))})
    monkeypatch.setattr(hostmem, "pci_address", lambda device: addr)
    assert hostmem.gpu_numa_node(0, sysfs) == 1
    assert hostmem.describe(0, sysfs) == {"pci": addr, "node": 1, "cpus": len(half), "bound": True, "source": "sysfs"}
    with hostmem.numa_local(0, sysfs) as node:
        assert node == 1 and os.sched_getaffinity(0) == set(half)
    assert os.sched_getaffinity(0) == set(mine)
    with pytest.raises(ValueError):               # restored when the body raises as well
        with hostmem.numa_local(0, sysfs):
            raise ValueError("body")
    assert os.sched_getaffinity(0) == set(mine)


@pytest.mark.parametrize("node,lists", [(-1, {0: "0-3"}), (0, {0: "100000-100003"}), (2, {0: "0-3"})])
def test_no_answer_is_a_no_op(tmp_path, monkeypatch, node, lists):
    mine = os.sched_getaffinity(0)
    addr = "0000:1b:00.0"
    sysfs = _fake_sysfs(tmp_path, addr, node, lists)
    monkeypatch.setattr(hostmem, "pci_address", lambda device: addr)
    monkeypatch.setattr(hostmem, "nvml_cpus", lambda device: set())       # no driver on this box
    assert hostmem.describe(0, sysfs)["bound"] is False
    with hostmem.numa_local(0, sysfs) as got:
        assert got is None and os.sched_getaffinity(0) == mine
    monkeypatch.setattr(hostmem, "pci_address", lambda device: None)       # no GPU / no PCI address
    assert hostmem.gpu_numa_node(0, sysfs) is None


def test_nvml_answer_is_used_when_sysfs_has_none(tmp_path, monkeypatch):
    mine = sorted(os.sched_getaffinity(0))
    if len(mine) < 2:
        pytest.skip("needs two CPUs in the affinity mask")
    half = set(mine[: len(mine) // 2])
    addr = "0000:c0:00.0"
    sysfs = _fake_sysfs(tmp_path, addr, -1, {0: "0-3"})
    monkeypatch.setattr(hostmem, "pci_address", lambda device: addr)
    monkeypatch.setattr(hostmem, "nvml_cpus", lambda device: half)
    assert hostmem.describe(0, sysfs) == {"pci": addr, "node": None, "cpus": len(half), "bound": True, "source": "nvml"}
    with hostmem.numa_local(0, sysfs) as node:
        assert node == -1 and os.sched_getaffinity(0) == half
    assert os.sched_getaffinity(0) == set(mine)


@pytest.mark.parametrize("B", [0, 1, 1000, 1 << 20, (1 << 20) + 1, 3 << 19, 2 << 20, (7 << 18) + 5, 1 << 22, (1 << 22) + 12345, 10_000_000])
def test_host_chunk_plan_covers_the_batch(B):
    """field._host_chunks splits a host-fed batch into pipelined pieces (short first and last ones): every point exactly
    once, no empty piece."""
    from nerf_navigation_b200.field import _host_chunks
    plan = _host_chunks(B)
    assert sum(plan) == B and all(n > 0 for n in plan)
