#!/usr/bin/env python
"""ORACLE INFRASTRUCTURE -- generates ``tests/golden/kuka_*.npz`` by running the REAL reference.

Run in the build container (needs ``/root/reference``):  ``python -m oracle.make_golden``.

Each fixture holds, for one seeded AllegroKukaReorientation scenario, the complete input state
(``in/<key>``), and per step the actions (``s<i>/actions``), the raw random table (``s<i>/rnd``), the
state after the reference's own ``pre_physics_step`` (``s<i>/pre/<key>``, plus the int32 actor rows it
handed to ``set_actor_root_state_tensor_indexed`` / ``set_dof_state_tensor_indexed``) and after its
``post_physics_step`` + ``VecTask.step`` tail (``s<i>/post/<key>``).  The fixtures travel to machines
without the reference; ``tests/test_oracle_golden.py`` (CPU) pins the restatement to them and
``tests/test_kuka_golden_gpu.py`` pins the CUDA path to them.
"""
from __future__ import annotations

import os
import sys
from typing import Dict

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import kuka_oracle as ko, kuka_state as ks, reference_loader as rl  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# name -> (n_envs, seed, fresh, steps, KukaConsts overrides)
SCENARIOS = {
    "kuka_reorient_fresh": (32, 101, True, 3, {}),
    "kuka_reorient_mid": (40, 202, False, 3, {}),
    "kuka_reorient_hold30_noforce": (24, 303, False, 2, dict(success_steps=30, max_episode_length=300, force_scale=0.0)),
    "kuka_reorient_tighter_tol": (24, 404, False, 2, dict(success_tolerance=0.04)),
    # the sibling subtasks of the single-arm task (allegro_kuka_regrasping.py, allegro_kuka_throw.py; env/regrasping.yaml, env/throw.yaml)
    "kuka_regrasp_fresh": (32, 505, True, 2, dict(subtask="regrasping", success_steps=30, max_episode_length=300)),
    "kuka_regrasp_mid": (40, 606, False, 3, dict(subtask="regrasping", success_steps=30, max_episode_length=300)),
    "kuka_throw_mid": (40, 707, False, 3, dict(subtask="throw", success_steps=5, max_episode_length=300, force_scale=0.0,
                                               success_tolerance=0.075, initial_tolerance=0.075, target_tolerance=0.075)),
}

# rigid_body_state is read-only on this path: only the input copy is stored
STATE_KEYS = ["root_state", "dof_state", "goal_states", "obs_buf", "rew_buf", "reset_buf", "reset_goal_buf",
              "progress_buf", "randomize_buf", "timeout_buf", "successes", "prev_episode_successes", "true_objective",
              "prev_episode_true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
              "closest_fingertip_dist", "furthest_hand_dist", "prev_targets", "cur_targets", "actions", "rb_forces",
              "random_force_prob", "rewards_episode", "episode_cumulative"]


def _snap(prefix: str, st: Dict[str, torch.Tensor], out: Dict[str, np.ndarray]) -> None:
    for k in STATE_KEYS:
        if k in st:
            out[f"{prefix}/{k}"] = st[k].detach().cpu().numpy().copy()


def _gym_rows(env, name: str) -> np.ndarray:
    """Concatenated int32 index arguments of every recorded ``gym.<name>`` call, and clears them."""
    rows = [args[2].to(torch.int32) for fn, args in env.gym.calls if fn == name]
    return torch.cat(rows).numpy().copy() if rows else np.zeros(0, dtype=np.int32)


def generate(name: str) -> str:
    n, seed, fresh, steps, over = SCENARIOS[name]
    c = ko.KukaConsts(**over)
    st = ks.make_state(n, seed=seed, c=c, fresh=fresh)
    out: Dict[str, np.ndarray] = {"meta/n": np.int64(n), "meta/seed": np.int64(seed), "meta/fresh": np.int64(fresh),
                                  "meta/steps": np.int64(steps)}
    for k, v in over.items():
        out[f"consts/{k}"] = np.str_(v) if isinstance(v, str) else np.float64(v)
    for k, v in st.items():
        out[f"in/{k}"] = v.numpy().copy()
    env = rl.make_reference_kuka(st, c)
    for i in range(steps):
        act, rnd = ks.make_actions(n, seed + i, c), ks.make_rnd(n, seed + i, c)
        out[f"s{i}/actions"], out[f"s{i}/rnd"] = act.numpy().copy(), rnd.numpy().copy()
        env.gym.calls.clear()
        with rl.scripted_randoms(rl.kuka_random_queue(st, c, rnd)) as q:
            env.pre_physics_step(act)
            assert len(q) == 0
        _snap(f"s{i}/pre", rl.harvest(env, st), out)
        out[f"s{i}/pre/root_rows"] = _gym_rows(env, "set_actor_root_state_tensor_indexed")
        out[f"s{i}/pre/hand_rows"] = _gym_rows(env, "set_dof_state_tensor_indexed")
        env.post_physics_step()
        env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)      # vec_task.py:424
        hv = rl.harvest(env, st)
        _snap(f"s{i}/post", hv, out)
        for k in ("successes", "true_objective_mean", "true_objective_min", "true_objective_max"):
            out[f"s{i}/extras/{k}"] = np.float32(env.extras[k])
        # the reference rebinds some attributes: carry them into the dict the queue builder reads next step
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
    only = sys.argv[1:]                       # optional: names of the fixtures to (re)generate; default all
    for name in (only or SCENARIOS):
        p = generate(name)
        print(f"{name}: {os.path.getsize(p) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
