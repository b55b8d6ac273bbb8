#!/usr/bin/env python
"""How long does the GPU itself need for K_post's algorithmic traffic?  Device-to-device copies of 20 MB (20 MB read + 20 MB written,
K_post's 40 MB) between rotating buffers (inputs never L2-resident), event-timed inside a CUDA graph like bench.py times the kernels.
Gives the practical ceiling of `roofline.frac` for a launch of this size (launch + ramp + drain are not free at 6 us of traffic)."""
import json
import torch

dev = "cuda:0"
out = {}
for mb in (5, 10, 20, 40, 80, 160):
    n = mb * 1000 * 1000 // 4
    pairs = max(2, 640 // mb)
    src = [torch.randn(n, device=dev) for _ in range(pairs)]
    dst = [torch.empty(n, device=dev) for _ in range(pairs)]
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        for a, b in zip(src, dst):
            b.copy_(a)
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for a, b in zip(src, dst):
                b.copy_(a)
        for _ in range(3):
            g.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 20
        e0.record(stream)
        for _ in range(reps):
            g.replay()
        e1.record(stream)
        stream.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * pairs)
    out[f"{mb}MB_read+{mb}MB_write"] = {"us": round(us, 2), "GB/s": round(2 * mb * 1e6 / (us * 1e-6) / 1e9, 1)}
print(json.dumps(out))
