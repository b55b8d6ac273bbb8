#!/usr/bin/env python
"""Step one of the side configs (bench.OTHER_CONFIGS key) a few times -- the target of ncu captures.
usage: python tools/profile_task.py cfg5_dextreme_adr [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

key = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 6
tasks, actions, post_b, pre_b = bench.build_other(key, "cuda:0", 2, 2)
for i in range(steps):
    bench.one_step(tasks[i % 2], actions[i % 2])
torch.cuda.synchronize()
print(key, "ok", post_b, pre_b)
