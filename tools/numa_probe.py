#!/usr/bin/env python
"""Host <-> device copy rates of pinned memory by the NUMA node the pages sit on (first touch under a CPU affinity).
   python tools/numa_probe.py"""
import glob, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
def cpulist(s):
    out = []
    for part in s.strip().split(","):
        if not part: continue
        a, _, b = part.partition("-")
        out += list(range(int(a), int(b or a) + 1))
    return out
nodes = {}
for d in sorted(glob.glob("/sys/devices/system/node/node[0-9]*")):
    nodes[int(d.rsplit("node", 1)[1])] = cpulist(open(d + "/cpulist").read())
print("nodes:", {k: (v[0], v[-1], len(v)) for k, v in nodes.items()})
prop = torch.cuda.get_device_properties(0)
bdf = "%04x:%02x:%02x.0" % (prop.pci_domain_id, prop.pci_bus_id, prop.pci_device_id) if hasattr(prop, "pci_bus_id") else None
print("gpu0 pci:", bdf)
if bdf:
    for f in ("numa_node", "local_cpulist"):
        p = f"/sys/bus/pci/devices/{bdf}/{f}"
        if os.path.exists(p): print(f, open(p).read().strip())
print("affinity now:", len(os.sched_getaffinity(0)), "cpus")
n = 1 << 30
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
dev2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
all_cpus = os.sched_getaffinity(0)
for node, cpus in list(nodes.items()) + [(-1, sorted(all_cpus))]:
    use = set(cpus) & all_cpus
    if not use: continue
    os.sched_setaffinity(0, use)
    a = torch.zeros(n, dtype=torch.uint8).pin_memory()     # first touch under this affinity
    b = torch.zeros(n, dtype=torch.uint8).pin_memory()
    res = {}
    for name, fn in (("h2d", lambda: dev.copy_(a, non_blocking=True)), ("d2h", lambda: b.copy_(dev2, non_blocking=True))):
        fn(); torch.cuda.synchronize()
        t = time.perf_counter()
        for _ in range(5): fn()
        torch.cuda.synchronize()
        res[name] = round(5 * n / (time.perf_counter() - t) / 1e9, 1)
    with torch.cuda.stream(s1): dev.copy_(a, non_blocking=True)
    with torch.cuda.stream(s2): b.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(5):
        with torch.cuda.stream(s1): dev.copy_(a, non_blocking=True)
        with torch.cuda.stream(s2): b.copy_(dev2, non_blocking=True)
    torch.cuda.synchronize()
    res["both_each_way"] = round(5 * n / (time.perf_counter() - t) / 1e9, 1)
    print("pages first touched on node", node, res, flush=True)
    del a, b
os.sched_setaffinity(0, all_cpus)
