#!/usr/bin/env python
"""ORACLE INFRASTRUCTURE -- generates ``tests/golden/hand_*.npz`` by running the REAL reference classes
(``ShadowHand`` / ``AllegroHand``).  Run in the build container: ``python -m oracle.make_golden_hand``.

Layout as in ``make_golden.py``: ``in/<key>`` the input state, per step ``s<i>/actions``, ``s<i>/rnd``,
``s<i>/pre/<key>`` (state after the reference's ``pre_physics_step`` + the actor rows it handed to the
simulator) and ``s<i>/post/<key>`` (after ``post_physics_step`` + the ``VecTask.step`` tail, including the
clamped ``obs_out`` / ``states_out`` that ``step()`` returns).  ``meta/consts_json`` holds the HandConsts.
"""
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

from oracle import hand_oracle as ho, hand_state as hs, reference_loader as rl  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# name -> (n_envs, seed, fresh, steps, HandConsts)
SCENARIOS = {
    "hand_shadow_full_state_fresh": (24, 11, True, 2, ho.HandConsts()),
    "hand_shadow_full_mid": (33, 12, False, 2, ho.HandConsts(obs_type="full", max_consecutive_successes=50, fall_penalty=-50.0)),
    "hand_shadow_openai_asym_forces": (24, 13, False, 2, ho.HandConsts(obs_type="openai", asymmetric=True, force_scale=1.0,
                                                                       act_moving_average=0.3)),
    "hand_allegro_full_state_mid": (40, 14, False, 3, ho.allegro_consts()),
    "hand_allegro_no_vel_relative": (24, 15, False, 2, ho.allegro_consts(obs_type="full_no_vel", use_relative_control=True,
                                                                          reset_dof_vel_noise=0.1, force_scale=2.0)),
}
STATE_KEYS = ["root_state", "dof_state", "goal_states", "obs_buf", "states_buf", "obs_out", "states_out", "rew_buf", "reset_buf",
              "reset_goal_buf", "progress_buf", "randomize_buf", "timeout_buf", "successes", "consecutive_successes", "prev_targets",
              "cur_targets", "actions", "rb_forces", "random_force_prob"]


def _snap(prefix, st, out):
    for k in STATE_KEYS:
        if k in st:
            out[f"{prefix}/{k}"] = st[k].detach().cpu().numpy().copy()


def _gym_rows(env, name):
    rows = [args[2].to(torch.int32) for fn, args in env.gym.calls if fn == name]
    return torch.cat(rows).numpy().copy() if rows else np.zeros(0, dtype=np.int32)


def generate(name: str) -> str:
    n, seed, fresh, steps, c = SCENARIOS[name]
    st = hs.make_state(n, seed, c, fresh=fresh)
    out: Dict[str, np.ndarray] = {"meta/steps": np.int64(steps), "meta/consts_json": np.array(json.dumps(dataclasses.asdict(c)))}
    for k, v in st.items():
        out[f"in/{k}"] = v.numpy().copy()
    env = rl.make_reference_hand(st, c)
    for i in range(steps):
        act, rnd = hs.make_actions(n, seed + i, c), hs.make_rnd(n, seed + i, c)
        out[f"s{i}/actions"], out[f"s{i}/rnd"] = act.numpy().copy(), rnd.numpy().copy()
        env.gym.calls.clear()
        with rl.scripted_randoms(rl.hand_random_queue(st, c, rnd)) as q:
            env.pre_physics_step(act)
            assert len(q) == 0
        hv = rl.harvest_hand(env, st, c)
        hv.pop("obs_out", None); hv.pop("states_out", None)
        _snap(f"s{i}/pre", hv, out)
        out[f"s{i}/pre/root_rows"] = np.sort(_gym_rows(env, "set_actor_root_state_tensor_indexed"))
        out[f"s{i}/pre/hand_rows"] = _gym_rows(env, "set_dof_state_tensor_indexed")
        env.post_physics_step()
        env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)      # vec_task.py:424
        hv = rl.harvest_hand(env, st, c)
        _snap(f"s{i}/post", hv, out)
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
    for name in SCENARIOS:
        p = generate(name)
        print(f"{name}: {os.path.getsize(p) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
