#!/usr/bin/env python3
"""D2H bandwidth of pinned host memory allocated under different CPU affinities (NUMA placement of the pinned pages).
The end-to-end number of bench.py is bound by the D2H copy of the CSR values; this tells how much of it is placement."""
import glob
import os
import subprocess
import time

import torch


def cpus_of(node):
    txt = open("/sys/devices/system/node/node%d/cpulist" % node).read().strip()
    out = set()
    for part in txt.split(","):
        if "-" in part:
            a, b = part.split("-")
            out.update(range(int(a), int(b) + 1))
        elif part:
            out.add(int(part))
    return out


def main():
    torch.cuda.set_device(0)
    nodes = sorted(int(p.rsplit("node", 1)[1]) for p in glob.glob("/sys/devices/system/node/node[0-9]*"))
    bus = torch.cuda.get_device_properties(0).pci_bus_id if hasattr(torch.cuda.get_device_properties(0), "pci_bus_id") else None
    try:
        q = subprocess.check_output(["nvidia-smi", "-i", "0", "--query-gpu=pci.bus_id", "--format=csv,noheader"], text=True).strip()
        bid = q.lower()[4:] if len(q) > 12 else q.lower()
        print("gpu0 bus", q, "numa_node", open("/sys/bus/pci/devices/%s/numa_node" % bid).read().strip())
    except Exception as e:
        print("numa_node lookup failed:", e, bus)
    print("nodes", nodes, "affinity now", len(os.sched_getaffinity(0)), "cpus")
    allowed = os.sched_getaffinity(0)
    n = 1 << 30
    d = torch.empty(n, dtype=torch.uint8, device="cuda")
    d.fill_(1)
    for node in nodes + [None]:
        if node is not None:
            c = cpus_of(node) & allowed
            if not c:
                print("node", node, "no allowed cpus")
                continue
            os.sched_setaffinity(0, c)
        else:
            os.sched_setaffinity(0, allowed)
        h = torch.empty(n, dtype=torch.uint8).pin_memory()
        h.fill_(0)
        torch.cuda.synchronize()
        for direction in ("d2h", "h2d"):
            ts = []
            for _ in range(5):
                t0 = time.perf_counter()
                if direction == "d2h":
                    h.copy_(d, non_blocking=True)
                else:
                    d.copy_(h, non_blocking=True)
                torch.cuda.synchronize()
                ts.append(time.perf_counter() - t0)
            print("pinned on node", node, direction, "%.1f GB/s" % (n / min(ts) / 1e9))
        del h
    os.sched_setaffinity(0, allowed)


if __name__ == "__main__":
    main()
