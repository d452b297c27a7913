"""utils/numa.py against a fake sysfs tree (no GPU needed for the parsing / lookup logic)."""
import os

from collective_communication_mpi_b200.utils import numa


def test_cpulist_parsing():
    assert numa.parse_cpulist("0-3,8,10-11\n") == [0, 1, 2, 3, 8, 10, 11]
    assert numa.parse_cpulist("5") == [5]
    assert numa.parse_cpulist("") == []


def test_lookup_in_a_fake_sysfs(tmp_path):
    addr = numa.pci_address(0, 0x1b, 0)
    assert addr == "0000:1b:00.0"
    dev = tmp_path / "bus" / "pci" / "devices" / addr
    dev.mkdir(parents=True)
    (dev / "numa_node").write_text("1\n")
    node = tmp_path / "devices" / "system" / "node" / "node1"
    node.mkdir(parents=True)
    (node / "cpulist").write_text("56-111\n")
    assert numa.node_of_pci_device(addr, str(tmp_path)) == 1
    assert numa.cpus_of_node(1, str(tmp_path)) == list(range(56, 112))
    # unknown device, and the "-1 = no affinity" answer virtual machines give
    assert numa.node_of_pci_device("0000:ff:00.0", str(tmp_path)) is None
    (dev / "numa_node").write_text("-1\n")
    assert numa.node_of_pci_device(addr, str(tmp_path)) is None
    assert numa.cpus_of_node(7, str(tmp_path)) == []


def test_binding_without_a_gpu_changes_nothing():
    before = os.sched_getaffinity(0)
    assert numa.bind_to_gpu_node(0) is None
    assert os.sched_getaffinity(0) == before
