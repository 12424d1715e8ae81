"""Copy-only host-to-device measurement: every rank copies a pinned 2 GiB host block to its GPU 8 times, all ranks at
once, no kernels.  Shows whether the end-to-end numbers of bench.py at N > 1 are bounded by the host side (memory /
PCIe root complexes shared by the GPUs) rather than by anything the library does.
    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 r02_h2d_copy_only.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("gloo")
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
rates = [None] * world
if world > 1:
    dist.all_gather_object(rates, gbs)
else:
    rates = [gbs]
if rank == 0:
    print(json.dumps({"n_gpus": world, "h2d_gbs_per_gpu": [round(r, 1) for r in rates], "h2d_gbs_total": round(sum(rates), 1)}))
if world > 1:
    dist.destroy_process_group()
