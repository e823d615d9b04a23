"""dist.bind_host_to_gpu: the rank's CPU affinity follows the NUMA node of its GPU (sysfs), and is left alone when the host exposes none."""
import os
import types

import pytest
import torch

from legged_tracking_b200 import dist


def _fake_sysfs(tmp_path, pci, node, cpulist):
    d = tmp_path / "bus/pci/devices" / pci
    d.mkdir(parents=True)
    (d / "numa_node").write_text(f"{node}\n")
    if node >= 0:
        n = tmp_path / "devices/system/node" / f"node{node}"
        n.mkdir(parents=True)
        (n / "cpulist").write_text(cpulist + "\n")
    return str(tmp_path)


@pytest.fixture
def restore_affinity():
    before = os.sched_getaffinity(0)
    yield before
    os.sched_setaffinity(0, before)


def test_cpulist_parser():
    assert dist._parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11}
    assert dist._parse_cpulist("\n") == set()


def test_binds_to_the_gpu_node(tmp_path, monkeypatch, restore_affinity):
    cpus = sorted(restore_affinity)
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: types.SimpleNamespace(pci_domain_id=0, pci_bus_id=0x1b, pci_device_id=0))
    root = _fake_sysfs(tmp_path, "0000:1b:00.0", 1, f"{cpus[0]},9999")
    info = dist.bind_host_to_gpu(0, sysfs=root)
    assert info == {"pci": "0000:1b:00.0", "numa_node": 1, "cpus": 1, "bound": True}
    assert os.sched_getaffinity(0) == {cpus[0]}


def test_no_numa_placement_changes_nothing(tmp_path, monkeypatch, restore_affinity):
    monkeypatch.setattr(torch.cuda, "get_device_properties", lambda i: types.SimpleNamespace(pci_domain_id=0, pci_bus_id=4, pci_device_id=0))
    info = dist.bind_host_to_gpu(0, sysfs=_fake_sysfs(tmp_path, "0000:04:00.0", -1, ""))
    assert info["numa_node"] == -1 and not info["bound"] and os.sched_getaffinity(0) == restore_affinity
    # a device sysfs does not list, or a node whose CPUs this process may not use: reported, not raised
    info = dist.bind_host_to_gpu(0, sysfs=str(tmp_path / "nowhere"))
    assert not info["bound"] and "error" in info and os.sched_getaffinity(0) == restore_affinity
