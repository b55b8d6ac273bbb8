#!/usr/bin/env python
"""Headline benchmark: task env-steps/s of the AllegroKuka reorientation hot path (BASELINE.json configs[2]).

One "step" = pre_physics_step + post_physics_step (+ the VecTask.step tail) over one block of
32768 envs per GPU, on synthetic simulator state (PhysX is the opaque producer).  Prints ONE JSON
line (see the driver contract in the task statement): ``value`` is device-resident throughput,
``e2e`` goes through ``VecTask.step`` with host buffers, ``roofline`` is the post-physics kernel
against the measured HBM peak, ``cpu_baseline`` is the CPU oracle port on this box's host cores.

``--impl reference`` times the reference's OWN CPU implementation of the same path instead of the CUDA path: the unmodified
``AllegroKukaReorientation.pre_physics_step`` / ``post_physics_step`` (allegro_kuka_base.py:1355-1443), imported from
``/root/reference`` or from the snapshot ``oracle/make_ref_snapshot.py`` leaves in ``oracle/_ref/`` at build time
(``cpu_baseline.kind`` "reference"); only if neither exists the oracle port is timed (``kind`` "port").

Timing method (whatever ``--steps`` / ``--warmup`` are): the K steps are whole CUDA-graph replays (a 16-step graph plus one
remainder graph, never an eager tail), at least 3 replays of every graph and one statistics all-reduce run before the first
event, and when K steps last less than 50 ms the K-step block is repeated ``reps`` times inside the timed region, each block
between its own pair of events; ``ms_per_step`` is the median block / K, ``timed_region_s`` the whole region.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

METRIC = "task env-steps/s (obs+reward+reset)"
UNIT = "env-steps/s"
WORKLOAD = "AllegroKuka reorientation (23-DOF, keypoint reward), 32768 envs/GPU"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=200)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--envs", type=int, default=32768, help="envs per GPU")
    ap.add_argument("--blocks", type=int, default=4, help="independent env blocks stepped round-robin so state does not stay in L2")
    ap.add_argument("--snapshots", type=int, default=4, help="rotating simulator snapshots per env block")
    ap.add_argument("--cpu-steps", type=int, default=0, help="steps of the CPU baseline sample (0 = auto, about 10-20 s)")
    ap.add_argument("--stats-interval", type=int, default=1, help="N > 1: graph replays between two all-reduces of the episode statistics")
    ap.add_argument("--stats", default="auto", choices=["auto", "peer", "nccl"],
                    help="N > 1: how the episode statistics cross the ranks.  peer: wait-free exchange over NVLink peer memory INSIDE the "
                         "post-physics kernel (HaKukaBuffers.xchg, no launch; checked against nccl at start-up); nccl: asynchronous "
                         "all_reduce once per graph replay; auto (default): peer if the symmetric-memory rendezvous and the check succeed")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--layout", default="both", choices=["both", "isaacgym", "soa"],
                    help="simulator state layout: component-major planes (HA_STATE_SOA, the north star's SoA state; headline), "
                         "Isaac Gym AoS rows, or both (default: headline on soa, the AoS numbers ride along)")
    ap.add_argument("--no-other-configs", action="store_true", help="skip the short runs of configs 1, 2, 4 and 5 that ride along at N=1")
    ap.add_argument("--l2-fetch", type=int, default=0, help="cudaLimitMaxL2FetchGranularity in bytes (0 = leave the driver default)")
    return ap.parse_args()


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


# ------------------------------------------------------------------------------------------------
# clocks
# ------------------------------------------------------------------------------------------------
class ClockSampler:
    """Samples SM clock and throttle reasons through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown"}

    def __init__(self, index: int):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None
        self.t = threading.Thread(target=self._run, daemon=True)

    def _run(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                mask = self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in self.REASONS.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(0.005)

    def __enter__(self):
        if self.nv is not None:
            self.t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self.nv is not None:
            self.t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


# ------------------------------------------------------------------------------------------------
# CPU arms
# ------------------------------------------------------------------------------------------------
def bench_config(args):
    """The workload description both arms print (identical keys and values, so the driver can match the two lines)."""
    B, S, N = args.blocks, args.snapshots, args.envs
    layout = "soa" if args.layout in ("both", "soa") else "isaacgym"
    return {"workload": WORKLOAD, "envs_per_gpu": N, "state_layout": layout, "env_blocks": B, "snapshots_per_block": S,
            "l2": f"inputs larger than L2: {B} env blocks x {S} simulator snapshots stepped round-robin (GPU arm; the CPU arm steps one block)",
            "launch": "GPU arm: CUDA graphs (16 steps + one remainder graph), whole replays only; CPU arm: eager torch on all host threads",
            "rng": "GPU arm: in-kernel Philox; CPU arm: torch's generator"}


def _cpu_loop(step_fn, n_envs, steps, warmup, budget_s):
    """Times ``steps`` calls of step_fn(i) after ``warmup`` untimed ones; ``steps <= 0`` or a sample that would exceed ``budget_s``
    is cut to a bounded sample (and the actual count is returned)."""
    step_fn(0)                                         # first call pays allocator / thread-pool / jit start-up
    t0 = time.perf_counter()
    step_fn(1)
    one = max(time.perf_counter() - t0, 1e-4)
    cap = max(3, int(budget_s / one))
    steps = cap if steps <= 0 else min(steps, cap)
    for i in range(max(0, warmup - 2)):
        step_fn(i)
    t0 = time.perf_counter()
    for i in range(steps):
        step_fn(i)
    dt = time.perf_counter() - t0
    return n_envs * steps / dt, steps, dt


def cpu_port_rate(n_envs: int, steps: int, warmup: int = 2, budget_s: float = 15.0):
    """env-steps/s of the CPU oracle port (torch eager on all host cores), pre + post per step."""
    from oracle import kuka_oracle as ko, kuka_state as ks
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = ko.KukaConsts()
    st = ks.make_state(n_envs, seed=1, c=c, fresh=False)
    acts = [ks.make_actions(n_envs, s) for s in range(2)]
    rnds = [ks.make_rnd(n_envs, s) for s in range(2)]

    def step(i):
        ko.pre_step(st, c, acts[i % 2], rnds[i % 2])
        ko.post_step(st, c)

    rate, steps, dt = _cpu_loop(step, n_envs, steps, warmup, budget_s)
    return rate, cores, steps, dt


def reference_available():
    try:
        from oracle import reference_loader as rl
        return rl.reference_available()
    except Exception:      # noqa: BLE001
        return False


def cpu_reference_rate(n_envs: int, steps: int, warmup: int = 2, budget_s: float = 15.0, device: str = "cpu"):
    """env-steps/s of the reference's own classes: AllegroKukaReorientation.pre_physics_step + post_physics_step
    (allegro_kuka_base.py:1355-1443, unmodified, torch's own generator), eager torch on all host cores (or on ``device``)."""
    from oracle import kuka_oracle as ko, kuka_state as ks, reference_loader as rl
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    c = ko.KukaConsts()
    st = ks.make_state(n_envs, seed=1, c=c, fresh=False)
    env = rl.make_reference_kuka(st, c, device=device)
    acts = [ks.make_actions(n_envs, s).to(device) for s in range(2)]
    torch.manual_seed(1234)

    def step(i):
        env.pre_physics_step(acts[i % 2])
        env.post_physics_step()
        env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)      # vec_task.py:424
        if device != "cpu":
            torch.cuda.synchronize()

    rate, steps, dt = _cpu_loop(step, n_envs, steps, warmup, budget_s)
    return rate, cores, steps, dt


def cpu_arm(n_envs, steps, warmup, budget_s):
    """(cpu_baseline dict, port dict or None): the reference's own code when it is reachable, else the oracle port."""
    if reference_available():
        rate, cores, k, dt = cpu_reference_rate(n_envs, steps, warmup, budget_s)
        cpu = {"value": rate, "unit": UNIT, "cores": cores, "kind": "reference", "steps": k,
               "sample": f"{k} steps x {n_envs} envs, the reference's own AllegroKukaReorientation.pre_physics_step + post_physics_step "
                         f"(allegro_kuka_base.py:1355-1443, unmodified) on CPU torch, {cores} threads, {dt:.1f} s"}
        return cpu
    rate, cores, k, dt = cpu_port_rate(n_envs, steps, warmup, budget_s)
    return {"value": rate, "unit": UNIT, "cores": cores, "kind": "port", "steps": k,
            "sample": f"{k} steps x {n_envs} envs, torch-CPU oracle port of pre+post physics step, {cores} threads, {dt:.1f} s "
                      "(reference snapshot oracle/_ref absent)"}


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    warmup = max(args.warmup, 3)
    cpu = cpu_arm(args.envs, args.steps, warmup, budget_s=90.0)
    steps, rate = cpu["steps"], cpu["value"]
    line = {
        "impl": "reference", "metric": METRIC, "value": rate, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": warmup, "ms_per_step": 1e3 * args.envs / rate, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic", "config": bench_config(args),
        "cpu_baseline": cpu,
        "e2e": {"value": rate, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if steps != args.steps:
        line["steps_timed"] = steps          # bounded sample: fewer steps than asked for were timed (CPU budget)
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
# CUDA arm
# ------------------------------------------------------------------------------------------------
def build_blocks(args, device, layout):
    import handarm_b200.config as hcfg
    from handarm_b200.sim import KUKA_LAYOUT, SyntheticSim
    from handarm_b200.tasks.allegro_kuka import AllegroKukaReorientation
    rank = int(os.environ.get("RANK", "0"))
    tasks, actions = [], []
    pad_kb = int(os.environ.get("HA_BENCH_PAD_KB", "0"))      # diagnostic: shifts every later allocation (the two timing states)
    if pad_kb:
        build_blocks.pad = torch.empty(pad_kb * 1024, dtype=torch.uint8, device=device)
    for b in range(args.blocks):
        cfg = hcfg.load_task_config("AllegroKuka", num_envs=args.envs, sim_device=device)
        sim = SyntheticSim(args.envs, KUKA_LAYOUT, device, seed=1234 + 100 * rank + b, snapshots=args.snapshots, state_layout=layout)
        task = AllegroKukaReorientation(cfg, device, device, 0, True, sim=sim, seed=1234 + rank)
        task.seed_rng(1234 + 100 * rank + b)
        sim.fill_kuka(goal_states=None)
        sim.record_calls = False
        gen = torch.Generator().manual_seed(99 + b + 100 * rank)
        actions.append((2.0 * torch.rand(args.envs, 23, generator=gen) - 1.0).to(device))
        tasks.append(task)
    return tasks, actions


# ---- the other configs of BASELINE.json (1 ShadowHand, 2 AllegroHand, 4 two-arm regrasping, 5 DeXtreme ADR): short runs that ride
# along with the headline at N=1, same method (CUDA graph over round-robin env blocks, device events, per-kernel graphs) ----
OTHER_CONFIGS = {
    "cfg1_shadow_hand": dict(task="ShadowHand", envs=4096, actions=20),
    "cfg2_allegro_hand": dict(task="AllegroHand", envs=16384, actions=16),
    "cfg4_two_arms_regrasping": dict(task="AllegroKukaTwoArmsLSTM", envs=8192, actions=46),
    "cfg5_dextreme_adr": dict(task="AllegroHandDextremeADR", envs=16384, actions=16),
}


def build_other(key, device, blocks, snapshots):
    import ctypes
    import handarm_b200.config as hcfg
    from handarm_b200 import sim as hsim
    from handarm_b200.tasks import allegro_kuka, dextreme, hand
    spec = OTHER_CONFIGS[key]
    n = spec["envs"]
    tasks, actions = [], []
    for b in range(blocks):
        seed = 4321 + b
        if key == "cfg4_two_arms_regrasping":
            cfg = hcfg.load_task_config(spec["task"], num_envs=n, sim_device=device, subtask="regrasping")
            sim = hsim.SyntheticSim(n, hsim.KUKA2_LAYOUT, device, seed=seed, snapshots=snapshots)
            task = allegro_kuka.AllegroKukaTwoArmsRegrasping(cfg, device, device, 0, True, sim=sim, seed=seed)
            sim.fill_kuka(goal_states=None)
        elif key == "cfg5_dextreme_adr":
            cfg = hcfg.load_task_config(spec["task"], num_envs=n, sim_device=device)
            sim = hsim.SyntheticSim(n, dextreme.DEXTREME_LAYOUT, device, seed=seed, snapshots=snapshots)
            task = dextreme.AllegroHandDextremeADR(cfg, device, device, 0, True, sim=sim, seed=seed)      # create_sim fills the snapshots
            task.adr_update_on_step = False      # the ADR range update reads its queue summary on the host (one sync): not capturable in a graph
        else:
            shadow = key == "cfg1_shadow_hand"
            cfg = hcfg.load_task_config(spec["task"], num_envs=n, sim_device=device)
            layout = hsim.SHADOW_HAND_LAYOUT if shadow else hsim.ALLEGRO_HAND_LAYOUT
            sim = hsim.SyntheticSim(n, layout, device, seed=seed, snapshots=snapshots)
            task = (hand.ShadowHand if shadow else hand.AllegroHand)(cfg, device, device, 0, True, sim=sim, seed=seed)
            sim.fill_hand(task.object_init_state[0, 0:3].tolist(), task.goal_states, task.dof_lower, task.dof_upper)
        task.seed_rng(seed)
        sim.record_calls = False
        gen = torch.Generator().manual_seed(77 + b)
        actions.append((2.0 * torch.rand(n, spec["actions"], generator=gen) - 1.0).to(device))
        tasks.append(task)
    t0 = tasks[0]
    fam = "dex" if key == "cfg5_dextreme_adr" else ("kuka" if key == "cfg4_two_arms_regrasping" else "hand")
    post_b = int(getattr(t0._lib, f"ha_{fam}_post_bytes_per_env_step")(ctypes.byref(t0._params)))
    pre_b = int(getattr(t0._lib, f"ha_{fam}_pre_bytes_per_env_step")(ctypes.byref(t0._params)))
    return tasks, actions, post_b, pre_b


def measure_other(key, device, steps, blocks=4, snapshots=2):
    hbm_gbs, _ = peaks()
    tasks, actions, post_b, pre_b = build_other(key, device, blocks, snapshots)
    n = OTHER_CONFIGS[key]["envs"]
    stream = torch.cuda.Stream()
    has_rna = getattr(tasks[0], "rna_network", None) is not None
    for t in tasks:                       # the memory-bound task kernels are timed on their own; the RNA chain (tensor cores) rides along below
        if has_rna:
            t.run_rna = False
    with torch.cuda.stream(stream):
        for t, a in zip(tasks, actions):
            for _ in range(3):
                one_step(t, a)
        stream.synchronize()
        for t in tasks:
            if hasattr(t, "frame_since_restart"):
                t.frame_since_restart, t.last_curriculum_update = 0, 0

        def graph_of(which):
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                for _ in range(snapshots):
                    for t, a in zip(tasks, actions):
                        if which in ("step", "pre"):
                            t.pre_physics_step(a)
                        if which in ("step", "post"):
                            t.sim.simulate()
                            t.post_physics_step()
            return g

        per_graph = blocks * snapshots

        def time_graph(g, region_ms=40.0):
            """us per step: >= 3 warm replays, then as many replays as fill ``region_ms`` between one pair of events"""
            for _ in range(3):
                g.replay()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record(stream)
            g.replay()
            a1.record(stream)
            stream.synchronize()
            k = int(min(20000, max(4, region_ms / max(a0.elapsed_time(a1), 1e-3))))
            a0.record(stream)
            for _ in range(k):
                g.replay()
            a1.record(stream)
            stream.synchronize()
            time_graph.last_reps = k
            return a0.elapsed_time(a1) * 1e3 / (k * per_graph)

        step_us = time_graph(graph_of("step"))
        reps = time_graph.last_reps
        post_us = time_graph(graph_of("post"))
        pre_us = time_graph(graph_of("pre"))
        rna = None
        if has_rna:                       # DeXtreme default config: random_network_adversary.enable = True (allegro_hand_dextreme.py:864-879)
            for t in tasks:
                t.run_rna = True
            for t, a in zip(tasks, actions):
                one_step(t, a)
            stream.synchronize()
            rna_step_us = time_graph(graph_of("step"))
            net = tasks[0].rna_network
            rna_us = rna_step_us - step_us
            rna = {"ms_per_step_with_rna": rna_step_us * 1e-3, "value_with_rna": n / (rna_step_us * 1e-6), "rna_ms": rna_us * 1e-3,
                   "rna_gflop": net.flops_per_env() * n / 1e9, "rna_fp32_equiv_tflops": net.flops_per_env() * n / (rna_us * 1e-6) / 1e12,
                   "rna_bf16_mma_tflops": 3 * (net.flops_per_env() - 2 * net.in_dims * net.num_feats1) * n / (rna_us * 1e-6) / 1e12,
                   "rna_impl": "tcgen05 split-bf16 (3 MMAs per product), fused epilogues"}
    out_extra = {"rna": rna} if rna else {}
    return {**out_extra, "envs_per_gpu": n, "value": n / (step_us * 1e-6), "unit": UNIT, "ms_per_step": step_us * 1e-3, "steps": reps * per_graph,
            "bytes_per_env": post_b + pre_b, "step_frac": n * (post_b + pre_b) / (step_us * 1e-6) / 1e9 / hbm_gbs,
            "post_kernel_us": post_us, "post_frac": n * post_b / (post_us * 1e-6) / 1e9 / hbm_gbs, "pre_kernel_us": pre_us}


def one_step(task, act):
    task.pre_physics_step(act)
    task.sim.simulate()
    task.post_physics_step()


def run_ours(args):
    import torch.distributed as dist
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    device = f"cuda:{local}"
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device(device))
    if args.l2_fetch:
        import ctypes
        rt = ctypes.CDLL("libcudart.so.12")
        rc = rt.cudaDeviceSetLimit(ctypes.c_int(5), ctypes.c_size_t(args.l2_fetch))      # cudaLimitMaxL2FetchGranularity
        got = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(got), ctypes.c_int(5))
        print(f"l2 fetch granularity: set rc={rc}, now {got.value} B", file=sys.stderr)
    layouts = ["soa", "isaacgym"] if args.layout == "both" else [args.layout]
    line = measure(args, layouts[0], device, rank, world, local, headline=True)
    if len(layouts) > 1:
        import gc
        gc.collect()
        torch.cuda.empty_cache()
        other = measure(args, layouts[1], device, rank, world, local, headline=False)
        if rank == 0:
            line["isaacgym_layout"] = {k: other[k] for k in ("value", "ms_per_step")}
            line["isaacgym_layout"]["roofline"] = {k: other["roofline"][k] for k in ("achieved", "frac", "kernel_us", "pre_kernel_us", "step_frac")}
            line["isaacgym_layout"]["note"] = ("same kernels on Isaac Gym's AoS tensors (root [N*A,13], rigid-body [N*B,13]): every 52-byte row "
                                               "costs whole 128-byte DRAM lines, 1.43 KB fetched per env for 0.61 KB used")
    if rank == 0 and world == 1 and not args.no_other_configs:
        import gc
        line["other_configs"] = {}
        for key in OTHER_CONFIGS:
            gc.collect()
            torch.cuda.empty_cache()
            line["other_configs"][key] = measure_other(key, device, min(args.steps, 800))
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def ncu_traffic(layout):
    """dram__bytes_read.sum + dram__bytes_write.sum per K_post launch from the committed ncu --set full capture (profiles/traffic.json)."""
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")
    try:
        return json.load(open(path))["kuka_post_kernel"][layout]
    except (OSError, KeyError, ValueError):
        return None


def ncu_traffic_source():
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "profiles", "traffic.json")
    try:
        return json.load(open(path)).get("_source")
    except (OSError, ValueError):
        return None


def measure(args, layout, device, rank, world, local, headline=True):
    import torch.distributed as dist
    hbm_gbs, peak_src = peaks()
    tasks, actions = build_blocks(args, device, layout)
    B, S, N = args.blocks, args.snapshots, args.envs
    stream = torch.cuda.Stream()
    stats_sum = torch.zeros(16, device=device)

    # first pass: every env resets (reset_buf starts as ones), then a few eager steps
    with torch.cuda.stream(stream):
        for t, a in zip(tasks, actions):
            for _ in range(2):
                one_step(t, a)
    stream.synchronize()
    for t in tasks:                       # curriculum host logic must not fire inside graph capture
        t.frame_since_restart = 0
        t.last_curriculum_update = 0

    # whole-box episode statistics once per graph replay (the cadence rl_games logs at: one epoch = 16 steps).  Default: an
    # asynchronous nccl all_reduce on nccl's own stream; --peer-stats: the engine's exchange over NVLink peer memory (one
    # single-CTA launch at the end of the step graph; falls back to nccl if the rendezvous fails)
    xchg, xchg_note = None, "single rank"
    if world > 1:
        from handarm_b200.stats import PeerStatsExchange, _MAX_SLOTS
        xchg_note = "nccl all_reduce, async, once per graph replay"
        if args.stats in ("auto", "peer"):
            try:
                xchg = PeerStatsExchange.connect(device=device)
                probe = torch.arange(16, device=device, dtype=torch.float32) * (rank + 1) + 0.25
                xchg.exchange(probe)
                torch.cuda.synchronize()
                dist.barrier()                              # the exchange never waits: make sure everyone has published number 1
                got = xchg.exchange(probe).clone()          # ... so that the second call returns its reduction
                want, wmax = probe.clone(), probe.clone()
                dist.all_reduce(want)
                dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
                want[_MAX_SLOTS] = wmax[_MAX_SLOTS]
                torch.cuda.synchronize()
                if not torch.allclose(got, want, rtol=1e-6, atol=0) or xchg.status() != 0:
                    raise RuntimeError(f"peer exchange disagrees with nccl: {got.tolist()} vs {want.tolist()}")
                xchg_note = (f"in-kernel exchange (HaKukaBuffers.xchg): peer stores over NVLink from one warp of the post-physics kernel, once "
                             f"per {B * S} steps, no launch; protocol checked against nccl all_reduce at start-up")
            except Exception as e:       # noqa: BLE001 -- any rendezvous / IPC failure: fall back, and say so
                xchg, xchg_note = None, f"nccl all_reduce (peer exchange unavailable: {type(e).__name__}: {e})"[:240]
            ok = torch.tensor([1 if xchg is not None else 0], device=device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok) == 0 and xchg is not None:
                xchg, xchg_note = None, "nccl all_reduce (peer exchange failed on another rank)"
            if xchg is not None:
                xchg.attach(tasks[0], interval=S)      # before the graphs are captured: the exchange rides inside tasks[0]'s post kernel, once per B*S steps (rl_games' epoch)

    # CUDA graphs of whole steps (2 kernels each) walking the B env blocks round-robin; every simulate() moves a block to its next
    # snapshot.  The main graph holds B*S steps, the remainder graph steps % (B*S): K steps are always whole replays.  The
    # statistics exchange (peer mode) rides at the end of the main graph (every argument of the launch lives in device memory).
    per_graph = B * S
    n_graph, rest = divmod(args.steps, per_graph)

    def capture(n_steps, with_xchg):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for i in range(n_steps):
                one_step(tasks[i % B], actions[i % B])
        return g

    graph = capture(per_graph, True)
    graph_rest = capture(rest, False) if rest else None
    replay_count = [0]

    def stats_allreduce():
        if world > 1 and xchg is None:          # one tiny all-reduce (sum half; the max half is analogous)
            stats_sum.copy_(tasks[0].stats, non_blocking=True)
            dist.all_reduce(stats_sum, async_op=True)

    def run_block():
        """Exactly args.steps steps: n_graph replays of the main graph + one of the remainder graph."""
        for _ in range(n_graph):
            graph.replay()
            replay_count[0] += 1
            if replay_count[0] % max(1, args.stats_interval) == 0:
                stats_allreduce()
        if graph_rest is not None:
            graph_rest.replay()
            if n_graph == 0:                    # fewer steps than one main graph: the block is the statistics cadence
                stats_allreduce()
        for t in tasks:
            t.advance_frames(max(1, args.steps // B))

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    with torch.cuda.stream(stream):
        # warm-up: >= 3 replays of every graph and one statistics all-reduce, whatever --warmup says; then --warmup steps' worth
        for _ in range(3):
            graph.replay()
            if graph_rest is not None:
                graph_rest.replay()
        stats_allreduce()
        blocks_warm = max(1, -(-max(args.warmup, 3) // max(args.steps, 1)))
        for _ in range(blocks_warm):
            run_block()
        barrier()
        # size of the timed region: one block probed with events; short blocks are repeated until the region is >= 50 ms
        p0, p1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        p0.record(stream)
        run_block()
        p1.record(stream)
        barrier()
        probe_ms = max(p0.elapsed_time(p1), 1e-3)
        reps = int(min(5000, max(1, -(-60.0 // probe_ms))))
        if world > 1:
            treps = torch.tensor([reps], device=device)
            dist.all_reduce(treps, op=dist.ReduceOp.MAX)
            reps = int(treps)
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
        barrier()
        with ClockSampler(local) as clk:
            t_wall = time.perf_counter()
            ev[0].record(stream)
            for r in range(reps):
                run_block()
                ev[r + 1].record(stream)
            barrier()
            t_wall = time.perf_counter() - t_wall
        block_ms = sorted(ev[r].elapsed_time(ev[r + 1]) for r in range(reps))
        ms = block_ms[reps // 2]                                  # median K-step block
        region_ms = ev[0].elapsed_time(ev[reps])
    if world > 1:
        tms = torch.tensor([ms, region_ms], device=device)
        dist.all_reduce(tms, op=dist.ReduceOp.MAX)
        ms, region_ms = float(tms[0]), float(tms[1])
    value = world * N * args.steps / (ms * 1e-3)
    value_mean = world * N * args.steps * reps / (region_ms * 1e-3)

    # ---- per-kernel device times: one CUDA graph of only pre launches, one of only post launches
    # (same round-robin over the env blocks, so inputs are as cold as in the timed region); CUDA
    # events on the launching stream around whole replays, divided by the launches per replay ----
    def kernel_graph(which):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for _ in range(S):
                for t, a in zip(tasks, actions):
                    if which == "pre":
                        t.pre_physics_step(a)
                    else:
                        t.sim.simulate()
                        t.post_physics_step()
        return g

    def time_graph(g, launches, region_ms=50.0):
        """us per launch: >= 3 warm replays, then as many replays as fill ``region_ms`` between one pair of events."""
        for _ in range(3):
            g.replay()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a0.record(stream)
        g.replay()
        a1.record(stream)
        stream.synchronize()
        k = int(min(20000, max(4, region_ms / max(a0.elapsed_time(a1), 1e-3))))
        a0.record(stream)
        for _ in range(k):
            g.replay()
        a1.record(stream)
        stream.synchronize()
        return a0.elapsed_time(a1) * 1e3 / (k * launches)

    with torch.cuda.stream(stream):
        g_post, g_pre = kernel_graph("post"), kernel_graph("pre")
        post_avg = time_graph(g_post, per_graph)
        pre_avg = time_graph(g_pre, per_graph)
    lib = tasks[0]._lib
    import ctypes
    post_bytes = int(lib.ha_kuka_post_bytes_per_env_step(ctypes.byref(tasks[0]._params)))
    pre_bytes = int(lib.ha_kuka_pre_bytes_per_env_step(ctypes.byref(tasks[0]._params)))
    achieved = N * post_bytes / (post_avg * 1e-6) / 1e9
    step_frac = (N * (post_bytes + pre_bytes) / (ms * 1e-3 / args.steps) / 1e9) / hbm_gbs

    # ---- L2-warm variant: a single env block stepped back to back (task state stays in L2) ----
    with torch.cuda.stream(stream):
        g1 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g1, stream=stream):
            for _ in range(S):
                one_step(tasks[0], actions[0])
        warm_value = N / (time_graph(g1, S) * 1e-6)

    # ---- concurrent env blocks: the B blocks are independent jobs, so a deployment that hosts several of them on one GPU may
    # step them on a stream each; the ramp and tail of one block's launches then overlap another block's (same graph size, same
    # cold inputs as the headline; reported beside it, the headline stays the serial step) ----
    conc_value = None
    if B > 1:
        with torch.cuda.stream(stream):
            side = [torch.cuda.Stream() for _ in tasks]
            gc = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gc, stream=stream):
                for s_, t, a in zip(side, tasks, actions):
                    s_.wait_stream(stream)
                    with torch.cuda.stream(s_):
                        for _ in range(S):
                            one_step(t, a)          # S steps of this block, back to back on its own stream
                for s_ in side:
                    stream.wait_stream(s_)
            conc_value = world * N / (time_graph(gc, per_graph) * 1e-6)

    # ---- end to end through VecTask.step with host buffers ----
    e2e = None
    if not args.no_e2e and headline:
        k = min(args.steps, 200)
        h_act = [(a.cpu()).pin_memory() for a in actions]
        t0 = tasks[0]
        h_obs = torch.empty((N, t0.num_obs), dtype=torch.float32).pin_memory()
        h_rew = torch.empty(N, dtype=torch.float32).pin_memory()
        h_rst = torch.empty(N, dtype=torch.int64).pin_memory()
        h_to = torch.empty(N, dtype=torch.bool).pin_memory()

        def e2e_step(i):
            t = tasks[i % B]
            obs, rew, rst, extras = t.step(h_act[i % B])          # H2D of the actions happens inside step()
            h_obs.copy_(obs["obs"], non_blocking=True)
            h_rew.copy_(rew, non_blocking=True)
            h_rst.copy_(rst, non_blocking=True)
            h_to.copy_(extras["time_outs"], non_blocking=True)
            torch.cuda.current_stream().synchronize()

        def timed_blocks(step_fn, warm):
            """median wall time of a k-step block (device idle on both sides of every block); blocks repeat until >= 0.3 s are timed"""
            for i in range(warm):
                step_fn(i)
            barrier()
            times, total = [], 0.0
            while total < 0.3 and len(times) < 200:
                t_start = time.perf_counter()
                for i in range(k):
                    step_fn(i)
                torch.cuda.synchronize()
                times.append(time.perf_counter() - t_start)
                total += times[-1]
                if world > 1:                   # every rank runs the same number of blocks
                    flag = torch.tensor([1.0 if (total < 0.3 and len(times) < 200) else 0.0], device=device)
                    dist.all_reduce(flag, op=dist.ReduceOp.MAX)
                    if float(flag) > 0:
                        total = min(total, 0.29)
                    else:
                        break
            barrier()
            return sorted(times)[len(times) // 2], len(times)

        with torch.cuda.stream(stream):
            dt_serial, _ = timed_blocks(e2e_step, 8)

        # the same calls with one stream and one set of pinned result buffers per env block: the copies of one block overlap the
        # kernels and copies of the others (PCIe is full duplex); a block's results are read on the host after its own stream drained
        streams = [torch.cuda.Stream() for _ in range(B)]
        h_out = [(torch.empty((N, t0.num_obs), dtype=torch.float32).pin_memory(), torch.empty(N, dtype=torch.float32).pin_memory(),
                  torch.empty(N, dtype=torch.int64).pin_memory(), torch.empty(N, dtype=torch.bool).pin_memory()) for _ in range(B)]

        def e2e_step_pipelined(i):
            blk = i % B
            streams[blk].synchronize()                             # the host owns this block's result buffers again
            with torch.cuda.stream(streams[blk]):
                obs, rew, rst, extras = tasks[blk].step(h_act[blk])
                h_out[blk][0].copy_(obs["obs"], non_blocking=True)
                h_out[blk][1].copy_(rew, non_blocking=True)
                h_out[blk][2].copy_(rst, non_blocking=True)
                h_out[blk][3].copy_(extras["time_outs"], non_blocking=True)

        stream.synchronize()
        dt, e2e_blocks = timed_blocks(e2e_step_pipelined, 2 * B)
        if world > 1:
            tdt = torch.tensor([dt, dt_serial], device=device)
            dist.all_reduce(tdt, op=dist.ReduceOp.MAX)
            dt, dt_serial = float(tdt[0]), float(tdt[1])
        v_multi, v_one = world * N * k / dt, world * N * k / dt_serial
        e2e = {"value": max(v_multi, v_one), "unit": UNIT, "h2d_bytes_per_step": N * 23 * 4,
               "d2h_bytes_per_step": N * (t0.num_obs * 4 + 4 + 8 + 1), "steps": k, "blocks_timed": e2e_blocks,
               "mode": ("VecTask.step per env block on its own stream" if v_multi >= v_one else "VecTask.step on one stream, host sync every step")
                       + f" ({B} blocks round-robin), pinned host actions in, obs / reward / reset / time_outs out every step; "
                         "median wall time of a block of `steps` calls",
               "value_stream_per_block": v_multi, "value_one_stream": v_one}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and headline:
        cpu = cpu_arm(N, args.cpu_steps, 3, budget_s=15.0)
        if cpu["kind"] == "reference":                       # the oracle port beside it (second figure), and the reference's eager
            rate, cores, csteps, cdt = cpu_port_rate(N, 0, budget_s=6.0)      # path on this GPU as a context number
            cpu["port"] = {"value": rate, "unit": UNIT, "cores": cores, "sample": f"{csteps} steps x {N} envs, torch-CPU oracle port, {cdt:.1f} s"}

    ref_cuda = None
    if rank == 0 and world == 1 and headline and not args.no_cpu_baseline and reference_available():
        try:
            rate, _, csteps, cdt = cpu_reference_rate(N, 0, 5, budget_s=4.0, device=device)
            ref_cuda = {"value": rate, "unit": UNIT, "ms_per_step": 1e3 * N / rate, "steps": csteps,
                        "what": "the reference's own classes stepped eagerly on this GPU (torch ATen launches, device-resident state, "
                                "host-synchronised every step): the 'before' of the production path, a context number"}
        except Exception as e:      # noqa: BLE001 -- context only
            ref_cuda = {"unavailable": f"{type(e).__name__}: {e}"[:200]}

    if rank == 0:
        cfg = bench_config(args)
        cfg["state_layout"] = layout
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": cfg,
            "reps": reps, "timed_region_s": region_ms * 1e-3, "value_mean_over_region": value_mean,
            "timing": f"median of {reps} blocks of {args.steps} steps, each between its own CUDA events on the launching stream; "
                      f"{n_graph} x {per_graph}-step graph + {rest}-step remainder graph per block; warm-up: 3 replays of every graph, "
                      f"one statistics all-reduce, {blocks_warm} block(s); max over ranks",
            "stats_exchange": xchg_note,
            "roofline": {"bound": "hbm", "kernel": f"kuka_post_kernel<K=4, SOA={int(layout == 'soa')}, NS=2>", "achieved": achieved, "peak": hbm_gbs, "unit": "GB/s",
                         "frac": achieved / hbm_gbs, "traffic": ncu_traffic(layout), "traffic_source": ncu_traffic_source(), "peak_source": peak_src, "kernel_us": post_avg,
                         "bytes_per_env": post_bytes, "pre_kernel_us": pre_avg, "pre_in_step_us": ms * 1e3 / args.steps - post_avg, "pre_bytes_per_env": pre_bytes,
                         "step_frac": step_frac, "kernel_timing": "CUDA events around >= 50 ms of graph replays of only-post / only-pre launches; the only-pre replays see no resets "
                                          "(the masks are consumed by the first launch), pre_in_step_us = step - post is the pre launch with the step's resets"},
            "value_l2_warm": warm_value,
            "value_concurrent_blocks": conc_value,
            "clocks": clk.summary(),
            "gpu_launches": 2 * args.steps * reps,
            "gpu_launches_per_step": 2,
        }
        if e2e is not None:
            line["e2e"] = e2e
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if ref_cuda is not None:
            line["reference_eager_cuda"] = ref_cuda
        return line
    return None


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
