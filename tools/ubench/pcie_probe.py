#!/usr/bin/env python
"""Host <-> device copy bandwidth of this box, per rank: pinned H2D, D2H and both at once, one rank at a time and all
ranks together, with the pinned buffer placed by default and after binding the process to the CPUs of the GPU's NUMA node.
Run alone (1 GPU) or under torchrun (N ranks). Prints one JSON line per rank 0 with every rank's numbers -- the
denominator of bench.py's end-to-end arm (e2e.pcie_frac)."""
import json
import os
import subprocess
import time

import torch
import torch.distributed as dist

GB = 1 << 30
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))


def barrier():
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def numa_of_gpu(i):
    try:
        p = torch.cuda.get_device_properties(i)
        bus = f"{p.pci_domain_id:04x}:{p.pci_bus_id:02x}:{p.pci_device_id:02x}.0"
        with open(f"/sys/bus/pci/devices/{bus}/numa_node") as f:
            return int(f.read()), bus
    except Exception as e:  # noqa: BLE001
        return None, str(e)


def cpus_of_node(n):
    try:
        with open(f"/sys/devices/system/node/node{n}/cpulist") as f:
            out = []
            for part in f.read().strip().split(","):
                a, _, b = part.partition("-")
                out += list(range(int(a), int(b or a) + 1))
            return out
    except Exception:  # noqa: BLE001
        return None


def measure(nbytes, reps=3):
    h_in = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_out = torch.empty(nbytes, dtype=torch.uint8, pin_memory=True)
    h_in.fill_(1), h_out.fill_(2)  # touch the pages here and now
    d_in = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    d_out = torch.ones(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def run(h2d, d2h):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(reps):
            if h2d:
                with torch.cuda.stream(s1):
                    d_in.copy_(h_in, non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2):
                    h_out.copy_(d_out, non_blocking=True)
        torch.cuda.synchronize()
        return nbytes * reps / (time.perf_counter() - t0) / 1e9

    run(True, True)
    return h_in, h_out, run


out = {"rank": rank, "gpu": local, "affinity_default": sorted(os.sched_getaffinity(0))[:4] + ["...", len(os.sched_getaffinity(0))]}
node, bus = numa_of_gpu(local)
out["gpu_pci"], out["gpu_numa_node"] = bus, node
res = {}
for placement in ("default", "numa_local"):
    if placement == "numa_local":
        cpus = cpus_of_node(node) if node is not None and node >= 0 else None
        if not cpus:
            res[placement] = "no NUMA information for this GPU (single node or virtualised): same as default"
            continue
        os.sched_setaffinity(0, set(cpus) & os.sched_getaffinity(0) or set(cpus))
    keep = measure(GB)
    run = keep[2]
    r = {}
    # one rank at a time
    for turn in range(world):
        barrier()
        if turn == rank:
            r["alone_h2d"], r["alone_d2h"] = round(run(True, False), 1), round(run(False, True), 1)
            both = run(True, True)
            r["alone_both_each"] = round(both, 1)
        barrier()
    # everybody together
    barrier()
    r["all_h2d"] = round(run(True, False), 1)
    barrier()
    r["all_d2h"] = round(run(False, True), 1)
    barrier()
    r["all_both_each"] = round(run(True, True), 1)
    barrier()
    res[placement] = r
    del keep
out["gbps"] = res
if world > 1:
    rows = [None] * world
    dist.all_gather_object(rows, out)
else:
    rows = [out]
if rank == 0:
    topo = ""
    try:
        topo = subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True, timeout=20).stdout
    except Exception:  # noqa: BLE001
        pass
    nodes = sorted(d for d in os.listdir("/sys/devices/system/node") if d.startswith("node")) if os.path.isdir("/sys/devices/system/node") else []
    print(json.dumps({"world": world, "host_numa_nodes": nodes, "ranks": rows, "note": "GB/s per rank; *_both_each = each direction while both run"}))
    print(topo)
if world > 1:
    dist.destroy_process_group()
