#!/usr/bin/env python
"""Time the random network adversary step (DeXtreme, 16384 envs by default) with CUDA events; prints one JSON line.
usage: python tools/rna_bench.py [n_envs] [reps]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from handarm_b200.rna import RandomNetworkAdversary  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 50
dev = "cuda:0"
net = RandomNetworkAdversary(n, 23, 16, 32, dev, seed=1)
dof_state = torch.randn(n, 16, 2, device=dev)
pose = torch.randn(n, 7, device=dev)
disc = torch.stack([torch.linspace(-0.5, 1.6, 32) for _ in range(16)]).to(dev)
for _ in range(5):
    net.perturbation(dof_state, pose, disc)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    net.perturbation(dof_state, pose, disc)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / reps
flops = net.flops_per_env() * n
print(json.dumps({"rna_envs": n, "ms_per_step": ms, "gflop_per_step": flops / 1e9, "tflops": flops / ms / 1e9,
                  "impl": getattr(net, "impl", "fp32 library GEMMs + hand-written ends")}))
