"""Diagnostic: device time of one ha_stats_exchange launch (a) with emulated ranks on one GPU, (b) between real ranks over symmetric
memory.  usage: python tools/xchg_bench.py   |   torchrun --nproc-per-node N tools/xchg_bench.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from handarm_b200.stats import PeerStatsExchange  # noqa: E402

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
dev = torch.device("cuda", torch.cuda.current_device())
v = torch.ones(16, device=dev)


def timed(fn, iters=200):
    for _ in range(10):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / iters


if world == 1:
    for w in (1, 2, 8):
        bufs = [torch.zeros(PeerStatsExchange.buffer_numel(w), device=dev) for _ in range(w)]
        rs = [PeerStatsExchange(r, w, bufs, dev) for r in range(w)]

        def all_ranks():
            for r in rs:
                r.exchange(v)
        print(f"emulated world {w}: {timed(all_ranks) / w:.2f} us per launch")
else:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=dev)
    x = PeerStatsExchange.connect(device=dev)
    t = timed(lambda: x.exchange(v))
    dist.barrier()
    s = torch.zeros(16, device=dev)
    t2 = timed(lambda: dist.all_reduce(s))
    print(f"rank {rank}/{world}: exchange {t:.2f} us per launch (back to back, lock-step with the peers), nccl all_reduce {t2:.2f} us, status {x.status()}")
    dist.destroy_process_group()
