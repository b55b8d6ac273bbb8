#!/usr/bin/env python
"""ORACLE INFRASTRUCTURE -- generates ``tests/golden/kuka2_*.npz`` by running the REAL reference class
``AllegroKukaTwoArmsRegrasping``.  Run in the build container: ``python -m oracle.make_golden_kuka2``.
Layout as in ``make_golden.py`` (``in/``, ``s<i>/actions``, ``s<i>/rnd``, ``s<i>/pre/``, ``s<i>/post/``, ``s<i>/extras/``)."""
from __future__ import annotations

import dataclasses
import json
import os
import sys
from typing import Dict

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import kuka2_oracle as k2, reference_loader as rl  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
SCENARIOS = {
    "kuka2_regrasp_fresh": (24, 21, True, 2, {}),
    "kuka2_regrasp_mid": (40, 22, False, 3, {}),
    "kuka2_regrasp_hold1_noforce_ema": (24, 23, False, 2, dict(success_steps=1, force_scale=0.0, act_moving_average=0.4)),
    "kuka2_reorient_fresh": (24, 24, True, 2, dict(subtask="reorientation")),
    "kuka2_reorient_mid": (40, 25, False, 3, dict(subtask="reorientation")),
}
STATE_KEYS = ["root_state", "dof_state", "goal_states", "obs_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf",
              "randomize_buf", "timeout_buf", "successes", "prev_episode_successes", "true_objective", "prev_episode_true_objective",
              "near_goal_steps", "lifted_object", "closest_keypoint_max_dist", "closest_fingertip_dist", "prev_targets", "cur_targets",
              "actions", "rb_forces", "random_force_prob", "rewards_episode", "episode_cumulative"]


def _snap(prefix, st, out):
    for k in STATE_KEYS:
        if k in st:
            out[f"{prefix}/{k}"] = st[k].detach().cpu().numpy().copy()


def _gym_rows(env, name):
    rows = [args[2].to(torch.int32) for fn, args in env.gym.calls if fn == name]
    return torch.cat(rows).numpy().copy() if rows else np.zeros(0, dtype=np.int32)


def generate(name: str) -> str:
    n, seed, fresh, steps, over = SCENARIOS[name]
    c = k2.Kuka2Consts(**over)
    st = k2.make_state(n, seed, c, fresh=fresh)
    out: Dict[str, np.ndarray] = {"meta/steps": np.int64(steps), "meta/consts_json": np.array(json.dumps(dataclasses.asdict(c)))}
    for k, v in st.items():
        out[f"in/{k}"] = v.numpy().copy()
    env = rl.make_reference_kuka2(st, c)
    for i in range(steps):
        act, rnd = k2.make_actions(n, seed + i, c), k2.make_rnd(n, seed + i, c)
        out[f"s{i}/actions"], out[f"s{i}/rnd"] = act.numpy().copy(), rnd.numpy().copy()
        env.gym.calls.clear()
        with rl.scripted_randoms(rl.kuka2_random_queue(st, c, rnd)) as q:
            env.pre_physics_step(act)
            assert len(q) == 0
        _snap(f"s{i}/pre", rl.harvest_kuka2(env, st), out)
        out[f"s{i}/pre/root_rows"] = _gym_rows(env, "set_actor_root_state_tensor_indexed")
        out[f"s{i}/pre/hand_rows"] = _gym_rows(env, "set_dof_state_tensor_indexed")
        env.post_physics_step()
        env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)
        hv = rl.harvest_kuka2(env, st)
        _snap(f"s{i}/post", hv, out)
        for k in ("successes", "true_objective_mean", "true_objective_min", "true_objective_max"):
            out[f"s{i}/extras/{k}"] = np.float32(env.extras[k])
        for k in ("reset_buf", "reset_goal_buf", "random_force_prob"):
            st[k] = hv[k]
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    np.savez_compressed(path, **out)
    return path


def main():
    if not rl.reference_available():
        raise SystemExit("the reference is not present here; golden vectors can only be generated in the build container")
    torch.manual_seed(0)
    import sys
    for name in (sys.argv[1:] or SCENARIOS):           # optional: only the named scenarios
        p = generate(name)
        print(f"{name}: {os.path.getsize(p) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
