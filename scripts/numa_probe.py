#!/usr/bin/env python
"""Where does osr_host_alloc put its pages?  For every visible GPU: the NUMA node of its PCIe root (sysfs), the nodes
of the pages of a 64 MiB osr_host_alloc block made with that GPU current (/proc/self/numa_maps), and the plain
pinned-copy rate from that block (CUDA events).  OSR_NO_NUMA_LOCAL=1 shows the placement without the library's help."""
import os
import re
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from puzzletunneldiscovery_b200 import pyosr  # noqa: E402


def block_nodes(addr):
    """{node: pages} of the mapping that starts at addr (numa_maps prints N<node>=<pages>)."""
    want = "%012x" % addr
    for line in open("/proc/self/numa_maps"):
        if line.split()[0].rjust(12, "0") == want:
            return {int(k): int(v) for k, v in re.findall(r"N(\d+)=(\d+)", line)}, line.strip()[:160]
    return {}, "(mapping not listed)"


def gpu_node(dev):
    bus = torch.cuda.get_device_properties(dev)
    pci = "%04x:%02x:%02x.0" % (getattr(bus, "pci_domain_id", 0), bus.pci_bus_id, bus.pci_device_id)
    try:
        return int(open("/sys/bus/pci/devices/%s/numa_node" % pci).read()), pci
    except OSError:
        return None, pci


def main():
    print("cpus allowed: %d; nodes: %s" % (len(os.sched_getaffinity(0)),
                                           sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node"))))
    for dev in range(torch.cuda.device_count()):
        torch.cuda.set_device(dev)
        torch.zeros(1, device="cuda")
        node, pci = gpu_node(dev)
        a = pyosr.pinned_empty(64 << 20, np.uint8)
        a[:] = 1
        nodes, line = block_nodes(a.ctypes.data)
        t = torch.from_numpy(a)
        d = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
        for _ in range(2):
            d.copy_(t, non_blocking=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            d.copy_(t, non_blocking=True)
        e1.record()
        torch.cuda.synchronize()
        gbs = 10 * (64 << 20) / (e0.elapsed_time(e1) * 1e-3) / 1e9
        print("gpu %d (%s) numa node %s: block pages by node %s; this thread on cpu %d; H2D %.1f GB/s"
              % (dev, pci, node, nodes, os.sched_getcpu() if hasattr(os, "sched_getcpu") else -1, gbs))
        print("   ", line)
        del t, a


if __name__ == "__main__":
    main()
