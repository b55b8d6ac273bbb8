"""Diagnostic (under torchrun): what bounds the end-to-end number at N ranks -- concurrent pinned D2H / H2D copies of the e2e step's
byte counts from every rank at once, next to the host's core count.  Not part of the product path.
usage: python -m torch.distributed.run --nproc-per-node N tools/pcie_diag.py"""
import json
import os
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
D2H, H2D, ITERS = 15761408, 3014656, 200
dev_out = torch.empty(D2H, dtype=torch.uint8, device="cuda")
host_out = torch.empty(D2H, dtype=torch.uint8).pin_memory()
dev_in = torch.empty(H2D, dtype=torch.uint8, device="cuda")
host_in = torch.empty(H2D, dtype=torch.uint8).pin_memory()


def barrier():
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def timed(fn):
    for _ in range(10):
        fn()
    barrier()
    t0 = time.perf_counter()
    for _ in range(ITERS):
        fn()
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    barrier()
    return dt / ITERS


s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def both():
    with torch.cuda.stream(s1):
        host_out.copy_(dev_out, non_blocking=True)
    with torch.cuda.stream(s2):
        dev_in.copy_(host_in, non_blocking=True)


res = {"rank": rank, "d2h_GBps": D2H / timed(lambda: host_out.copy_(dev_out, non_blocking=True)) / 1e9,
       "h2d_GBps": H2D / timed(lambda: dev_in.copy_(host_in, non_blocking=True)) / 1e9,
       "duplex_step_ms": timed(both) * 1e3}
t0 = time.perf_counter()
x = 0
for i in range(2_000_000):
    x += i
res["py_loop_s"] = time.perf_counter() - t0          # host core contention shows here
out = [None] * world
if world > 1:
    dist.all_gather_object(out, res)
else:
    out = [res]
if rank == 0:
    print(json.dumps({"world": world, "cpus": os.cpu_count(), "affinity": len(os.sched_getaffinity(0)),
                      "loadavg": os.getloadavg(), "ranks": out}))
if world > 1:
    dist.destroy_process_group()
