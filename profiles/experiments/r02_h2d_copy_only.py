"""Copy-only host-to-device measurement: every rank copies a pinned 2 GiB host block to its GPU 8 times, all ranks at
once, no kernels.  Shows whether the end-to-end numbers of bench.py at N > 1 are bounded by the host side (memory /
PCIe root complexes shared by the GPUs) rather than by anything the library does.  With numa_local the pinned block is
allocated while the process is confined to the CPUs next to its GPU (sysfs local_cpulist), so first touch puts it on
the GPU's own memory node.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 r02_h2d_copy_only.py [numa_local]"""
import json
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")


def parse_cpulist(text):
    cpus = set()
    for part in text.strip().split(","):
        if not part:
            continue
        a, _, b = part.partition("-")
        cpus.update(range(int(a), int(b or a) + 1))
    return cpus


p = torch.cuda.get_device_properties(local)
sysfs = "/sys/bus/pci/devices/%04x:%02x:%02x.0" % (p.pci_domain_id, p.pci_bus_id, p.pci_device_id)
node, cpus = None, set()
try:
    node = int(open(sysfs + "/numa_node").read())
    cpus = parse_cpulist(open(sysfs + "/local_cpulist").read())
except OSError:
    pass
before = os.sched_getaffinity(0)
if "numa_local" in sys.argv and cpus:
    os.sched_setaffinity(0, cpus & before or before)
n = 2 << 30
host = torch.empty(n, dtype=torch.uint8).pin_memory()
host.fill_(1)
dev = torch.empty(n, dtype=torch.uint8, device="cuda")
dev.copy_(host, non_blocking=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
for _ in range(8):
    dev.copy_(host, non_blocking=True)
torch.cuda.synchronize()
dt = time.perf_counter() - t0
gbs = 8 * n / dt / 1e9
info = {"gbs": round(gbs, 1), "numa_node": node, "local_cpus": len(cpus), "allowed_cpus": len(before)}
infos = [None] * world
if world > 1:
    dist.all_gather_object(infos, info)
else:
    infos = [info]
if rank == 0:
    print(json.dumps({"n_gpus": world, "numa_local_alloc": "numa_local" in sys.argv, "h2d_gbs_per_gpu": [i["gbs"] for i in infos],
                      "h2d_gbs_total": round(sum(i["gbs"] for i in infos), 1), "gpu_numa_node": [i["numa_node"] for i in infos],
                      "local_cpus": [i["local_cpus"] for i in infos], "allowed_cpus": infos[0]["allowed_cpus"]}))
if world > 1:
    dist.destroy_process_group()
