#!/usr/bin/env python
"""ORACLE INFRASTRUCTURE -- generates ``tests/golden/dextreme_*.npz`` by running the REAL reference class
``AllegroHandDextremeADR`` (``apply_randomizations`` stubbed, see ``reference_loader.make_reference_dextreme``).
Run in the build container: ``python -m oracle.make_golden_dextreme``.
Layout: ``in/``, ``s<i>/actions``, ``s<i>/rnd``, ``s<i>/prnd``, ``s<i>/pre/``, ``s<i>/post/``."""
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

from oracle import dextreme_oracle as dx, dextreme_state as ds, reference_loader as rl  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
SCENARIOS = {
    "dextreme_adr_fresh": (24, 31, 1, 3, {}),
    "dextreme_adr_mid": (40, 32, 0, 3, {}),
    "dextreme_adr_capped_axisrot_hold1": (24, 33, 0, 2, dict(use_capped_dof_control=True, apply_random_quat=False, num_success_hold_steps=1,
                                                            fall_penalty=-50.0, force_scale=0.0, act_moving_average=0.15)),
}
STATE_KEYS = ["rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "hold_count_buf", "apply_reset_buf", "successes",
              "consecutive_successes", "best_rotation_dist", "prev_targets", "cur_targets", "prev_prev_targets", "prev_actions",
              "prev_actions_queue", "rb_forces", "random_force_prob", "object_pose", "object_pose_wrt_wrist", "obs_object_pose",
              "obs_object_pose_freq", "timeout_buf", "actions", "dof_delta", "episode_cumulative", "obs_rows", "obs_out", "root_state",
              "dof_state", "goal_states", "frame"]


def _snap(prefix, st, out):
    for k in STATE_KEYS:
        if k in st:
            out[f"{prefix}/{k}"] = st[k].detach().cpu().numpy().copy()


def _gym_rows(env, name):
    rows = [args[2].to(torch.int32) for fn, args in env.gym.calls if fn == name]
    return torch.cat(rows).numpy().copy() if rows else np.zeros(0, dtype=np.int32)


def generate(name: str) -> str:
    n, seed, fresh, steps, over = SCENARIOS[name]
    c = dx.DexConsts(**over)
    st = ds.make_state(n, seed, c, fresh=fresh)
    out: Dict[str, np.ndarray] = {"meta/steps": np.int64(steps), "meta/consts_json": np.array(json.dumps(dataclasses.asdict(c)))}
    for k, v in st.items():
        out[f"in/{k}"] = v.numpy().copy()
    env = rl.make_reference_dextreme(st, c)
    for i in range(steps):
        act, rnd, prnd = ds.make_actions(n, seed + i, c), ds.make_rnd(n, seed + i, c), ds.make_prnd(n, seed + i, c)
        out[f"s{i}/actions"], out[f"s{i}/rnd"], out[f"s{i}/prnd"] = act.numpy().copy(), rnd.numpy().copy(), prnd.numpy().copy()
        env.gym.calls.clear()
        q_pre, q_post = rl.dextreme_random_queues(st, c, rnd, prnd)
        progress = torch.clamp((rnd[:, dx.RND_PROGRESS] * c.max_episode_length).long(), max=c.max_episode_length - 1)
        old_randint = torch.randint
        torch.randint = lambda *a, **k: progress.clone()
        try:
            with rl.scripted_randoms(q_pre) as q:
                env.pre_physics_step(act)
                assert len(q) == 0
            _snap(f"s{i}/pre", rl.harvest_dextreme(env, st, c), out)
            out[f"s{i}/pre/root_rows"] = _gym_rows(env, "set_actor_root_state_tensor_indexed")
            out[f"s{i}/pre/hand_rows"] = _gym_rows(env, "set_dof_state_tensor_indexed")
            with rl.scripted_randoms(q_post) as q:
                env.post_physics_step()
                assert len(q) == 0
        finally:
            torch.randint = old_randint
        env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)
        hv = rl.harvest_dextreme(env, st, c)
        _snap(f"s{i}/post", hv, out)
        for k in ("reset_buf", "reset_goal_buf", "apply_reset_buf", "random_force_prob", "progress_buf"):
            st[k] = hv[k]
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, name + ".npz")
    np.savez_compressed(path, **out)
    return path


def main():
    if not rl.reference_available():
        raise SystemExit("the reference is not present here; golden vectors can only be generated in the build container")
    torch.manual_seed(0)
    for name in SCENARIOS:
        p = generate(name)
        print(f"{name}: {os.path.getsize(p) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
