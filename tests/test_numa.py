"""NUMA placement of a rank (psb200/utils/numa.py), against a fake sysfs tree: which node a GPU hangs off, which CPUs
that node has, what a rank does with it — in a child process, so that the test runner keeps its own affinity."""
import json
import os
import subprocess
import sys

import pytest

from psb200.utils import numa

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _fake_sysfs(root, nodes, devices):
    for node, cpulist in nodes.items():
        d = root / "devices" / "system" / "node" / f"node{node}"
        d.mkdir(parents=True)
        (d / "cpulist").write_text(cpulist + "\n")
    (root / "devices" / "system" / "node" / "online").write_text("0-1\n")
    for bdf, node in devices.items():
        d = root / "bus" / "pci" / "devices" / bdf
        d.mkdir(parents=True)
        (d / "numa_node").write_text(f"{node}\n")
    return str(root)


def test_sysfs_parsing(tmp_path):
    assert numa.parse_cpulist("0-3,8,10-11\n") == {0, 1, 2, 3, 8, 10, 11} and numa.parse_cpulist("") == set()
    assert numa.pci_address(0, 0x1b, 0) == "0000:1b:00.0" and numa.pci_address(1, 0xc3, 2) == "0001:c3:02.0"
    sysfs = _fake_sysfs(tmp_path, {0: "0-55,112-167", 1: "56-111,168-223"}, {"0000:1b:00.0": 0, "0000:c3:00.0": 1,
                                                                              "0000:05:00.0": -1})
    assert numa.online_nodes(sysfs) == {0, 1}
    assert numa.node_of_pci_device("0000:c3:00.0", sysfs) == 1 and numa.node_of_pci_device("0000:1b:00.0", sysfs) == 0
    assert numa.node_of_pci_device("0000:05:00.0", sysfs) is None            # -1: the platform does not say
    assert numa.node_of_pci_device("0000:ff:00.0", sysfs) is None            # no such device
    assert len(numa.cpus_of_node(1, sysfs)) == 112 and numa.cpus_of_node(7, sysfs) == set()


def test_nothing_happens_without_a_second_node_or_when_switched_off(tmp_path, monkeypatch):
    one = _fake_sysfs(tmp_path / "one", {0: "0-7"}, {"0000:1b:00.0": 0})
    before = os.sched_getaffinity(0)
    assert numa.bind_to_node(0, one) is None and numa.bind_to_node(None, one) is None
    assert numa.bind_to_device_node(0, one) is None                         # no CUDA device here: swallowed
    monkeypatch.setenv("PSB_NUMA_BIND", "0")
    assert numa.bind_to_device_node(0) is None
    assert os.sched_getaffinity(0) == before


def test_a_rank_moves_to_the_cpus_of_its_node(tmp_path):
    allowed = sorted(os.sched_getaffinity(0))
    if len(allowed) < 2:
        pytest.skip("needs two CPUs")
    half = allowed[len(allowed) // 2:]
    sysfs = _fake_sysfs(tmp_path, {0: ",".join(map(str, allowed[: len(allowed) // 2])), 1: ",".join(map(str, half)),
                                   2: "4000-4003"}, {})
    code = ("import json, os, sys, threading; sys.path.insert(0, %r); from psb200.utils import numa\n"
            "r = numa.bind_to_node(1, %r); mine = sorted(os.sched_getaffinity(0)); seen = []\n"
            "t = threading.Thread(target=lambda: seen.append(sorted(os.sched_getaffinity(0)))); t.start(); t.join()\n"
            "far = numa.bind_to_node(2, %r)\n"
            "print(json.dumps({'r': r, 'mine': mine, 'thread': seen[0], 'far': far, 'after': sorted(os.sched_getaffinity(0))}))"
            % (ROOT, sysfs, sysfs))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    got = json.loads(out.stdout)
    assert got["r"]["node"] == 1 and got["r"]["cpus_bound"] and got["r"]["cpus"] == len(half)
    assert got["mine"] == half and got["thread"] == half                    # threads started afterwards inherit it
    # a node none of whose CPUs the process may use: the CPUs stay as they are (the memory policy is still tried)
    assert got["far"]["node"] == 2 and not got["far"]["cpus_bound"] and got["after"] == half


def test_memory_policy_call_is_well_formed():
    """On this machine node 0 exists: preferring it must be accepted by the kernel; a node that does not exist is
    refused without raising.  (Run in a child: the policy sticks to the thread.)"""
    code = ("import sys; sys.path.insert(0, %r); from psb200.utils import numa\n"
            "print(numa.prefer_memory_of(0), numa.prefer_memory_of(900), numa.prefer_memory_of(-1))" % ROOT)
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr
    if os.uname().machine in numa.SYS_SET_MEMPOLICY:
        assert out.stdout.split() == ["True", "False", "False"]
