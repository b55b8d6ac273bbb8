"""Diagnostic for the two timing states ("Step"): run the kernel-only bench after loading OTHER CUDA modules first, which
moves where the driver places this library's code.  usage (under gpurun): HA_PRELOAD=<n> python tools/preload_probe.py"""
import ctypes
import json
import os
import runpy
import shutil
import sys
import tempfile

import numpy as np
import torch

n = int(os.environ.get("HA_PRELOAD", "0"))
root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, root)
os.chdir(root)
if n:
    torch.zeros(1, device="cuda")
    src = os.path.join(root, "isaacgym-hand-arm_b200", "libhandarm_b200_trace.so")
    tmp = tempfile.mkdtemp()
    keep = []
    for i in range(n):                                    # distinct files -> distinct modules
        dst = os.path.join(tmp, f"pre{i}.so")
        shutil.copyfile(src, dst)
        lib = ctypes.CDLL(dst)
        buf = np.zeros(2 * 4096 * 136 // 8, dtype=np.int64)
        lib.ha_debug_trace_read(buf.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(buf.nbytes))      # forces the module load
        keep.append(lib)
sys.argv = ["bench.py", "--steps", "800", "--warmup", "50", "--no-e2e", "--no-cpu-baseline", "--no-other-configs", "--layout", "soa"]
import io
import contextlib
out = io.StringIO()
with contextlib.redirect_stdout(out):
    try:
        runpy.run_path(os.path.join(root, "bench.py"), run_name="__main__")
    except SystemExit:
        pass
d = json.loads(out.getvalue().strip().splitlines()[-1])
r = d["roofline"]
print("preload", n, "step_us", round(d["ms_per_step"] * 1e3, 2), "post_us", round(r["kernel_us"], 2), "pre_us", round(r["pre_kernel_us"], 2))
