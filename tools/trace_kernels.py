"""In-kernel timeline of the AllegroKuka step kernels (diagnostic build: make -C isaacgym-hand-arm_b200/csrc trace).

    HANDARM_B200_LIB=$PWD/isaacgym-hand-arm_b200/libhandarm_b200_trace.so python tools/trace_kernels.py [--envs N] [--layout soa]

Replays the bench's step graph (4 env blocks round-robin), then reads the per-CTA records of the LAST pre and post launch:
%globaltimer at CTA start / end (cross-SM alignment, coarse) and SM clock at named points (fine, per CTA).  Prints, per kernel,
when CTAs start relative to the first one, how long each phase takes (median / p90 / max over CTAs) and the span of the launch."""
import argparse
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--envs", type=int, default=32768)
ap.add_argument("--layout", default="soa")
ap.add_argument("--blocks", type=int, default=4)
ap.add_argument("--snapshots", type=int, default=4)
ap.add_argument("--which", default="step", choices=["step", "post", "pre"])
ap.add_argument("--ghz", type=float, default=1.965)
a = ap.parse_args()

dev = "cuda:0"
torch.cuda.set_device(0)
tasks, actions = bench.build_blocks(a, dev, a.layout)
lib = tasks[0]._lib
assert hasattr(lib, "ha_debug_trace_read"), "not a trace build: set HANDARM_B200_LIB to libhandarm_b200_trace.so"
stream = torch.cuda.Stream()
with torch.cuda.stream(stream):
    for t, x in zip(tasks, actions):
        for _ in range(2):
            bench.one_step(t, x)
    stream.synchronize()
    for t in tasks:
        t.frame_since_restart = 0
        t.last_curriculum_update = 0
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g, stream=stream):
        for i in range(a.blocks * a.snapshots):
            t, x = tasks[i % a.blocks], actions[i % a.blocks]
            if a.which in ("step", "pre"):
                t.pre_physics_step(x)
            if a.which in ("step", "post"):
                t.sim.simulate()
                t.post_physics_step()
    for _ in range(20):
        g.replay()
    stream.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(200):
        g.replay()
    e1.record(stream)
    stream.synchronize()
    print(f"{a.which}: {e0.elapsed_time(e1) * 1e3 / (200 * a.blocks * a.snapshots):.2f} us per graph step (trace build)")

MAXC = 4096
rec = np.dtype([("gt0", "<u8"), ("gt1", "<u8"), ("c", "<i8", (14,)), ("smid", "<i4"), ("n", "<i4")])
buf = np.zeros((2, MAXC), dtype=rec)
rc = lib.ha_debug_trace_read(buf.ctypes.data_as(ctypes.c_void_p), ctypes.c_size_t(buf.nbytes))
assert rc == 0


def pct(x):
    x = np.asarray(x, dtype=np.float64)
    return f"med {np.median(x):7.2f}  p90 {np.percentile(x, 90):7.2f}  max {x.max():7.2f}"


names = {0: ["start", "loads issued, masks in", "ballots", "reset w0: DOF rows stored", "reset w0: draws in hand", "barrier", "element pass stored", "reset w3: bookkeeping stored", "reset w3: lists written", "end",
             "reset w0: its envs done", "reset w0: rotation half stored", "reset w0: object_init_state arrived"],
         1: ["start", "prologue+issue", "t0 B1 inputs", "t0 role0 tips done", "t0 role1 keypoints done", "t0 role2 dofs done", "t0 role3 copies done",
             "t0 B2", "t0 reward chain done", "t0 B3", "t1 B1 inputs", "t1 B2", "t1 B3", "end"]}
spans = {}
for k, label in ((0, "kuka_pre_kernel"), (1, "kuka_post_kernel")):
    r = buf[k]
    r = r[r["n"] > 0]
    if len(r) == 0:
        continue
    g0 = r["gt0"].min()
    spans[k] = (g0, r["gt1"].max())
    print(f"\n== {label}: {len(r)} CTAs on {len(set(r['smid']))} SMs; launch span {(r['gt1'].max() - g0) / 1e3:.2f} us (globaltimer)")
    print(f"   CTA start after the first CTA [us]: {pct((r['gt0'] - g0) / 1e3)}")
    print(f"   CTA end   after the first CTA [us]: {pct((r['gt1'] - g0) / 1e3)}")
    print(f"   CTA lifetime (globaltimer) [us]    : {pct((r['gt1'] - r['gt0']) / 1e3)}")
    cats = [("all", np.ones(len(r), bool))]
    if k == 0:
        kind, n_reset = r["c"][:, 13] % 16, r["c"][:, 13] // 16
        cats += [("no reset", kind == 0), ("goal reset only", kind == 1), ("full reset", kind == 2),
                 ("1 env with a reset", n_reset == 1), ("2 envs", n_reset == 2), ("3+ envs", n_reset >= 3)]
    else:
        ntile = np.where(r["c"][:, 10] > 0, 2, 1)
        cats += [("1 tile", ntile == 1), ("2+ tiles", ntile == 2)]
    for cname, m in cats:
        if m.sum() == 0:
            continue
        print(f"   -- {cname}: {int(m.sum())} CTAs; time since CTA start [us at {a.ghz} GHz]; last warp's end (globaltimer) {pct((r['gt1'][m] - r['gt0'][m]) / 1e3)}")
        for i, nm in enumerate(names[k]):
            idx = i
            v = r["c"][m][:, idx]
            ok = v > 0
            if i == 0 or ok.sum() == 0:
                continue
            d = (v[ok] - r["c"][m][ok, 0]) / (a.ghz * 1e3)
            print(f"      {nm:22s} {pct(d)}   ({int(ok.sum())} CTAs)")
if 0 in spans and 1 in spans:
    print(f"\npre end -> post first CTA start: {(int(spans[1][0]) - int(spans[0][1])) / 1e3:.2f} us;  pre start -> post end: {(int(spans[1][1]) - int(spans[0][0])) / 1e3:.2f} us")
