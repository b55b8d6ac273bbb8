"""ORACLE INFRASTRUCTURE (tests / bench CPU baseline only) -- seeded ShadowHand / AllegroHand test states.

Same role as ``kuka_state.py``: a dict of CPU tensors (simulator tensors in Isaac Gym's layout + every
per-env task tensor) that can be injected into the reference classes, fed to the restatement
(``hand_oracle``) and loaded into the CUDA task (``tests/hand_harness.py``).
"""
from __future__ import annotations

from typing import Dict

import torch

from .hand_oracle import HandConsts, rnd_stride, RND_FORCE_NORMAL

# joint limits: assets/mjcf/open_ai_assets/hand/robot.xml (24 joints, WRJ1 WRJ0 FFJ3..0 MFJ3..0 RFJ3..0 LFJ4..0 THJ4..0)
SHADOW_LOWER = [-0.489, -0.698, -0.349, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0,
                -1.047, 0.0, -0.209, -0.524, -1.571]
SHADOW_UPPER = [0.14, 0.489, 0.349, 1.571, 1.571, 1.571, 0.349, 1.571, 1.571, 1.571, 0.349, 1.571, 1.571, 1.571, 0.785, 0.349, 1.571,
                1.571, 1.571, 1.047, 1.222, 0.209, 0.524, 0.0]
# assets/urdf/kuka_allegro_description/allegro_touch_sensor.urdf (16 joints)
ALLEGRO_LOWER = [-0.558488888889, -0.279244444444, -0.279244444444, -0.279244444444] * 3 + [0.279244444444, -0.331602777778,
                                                                                         -0.279244444444, -0.279244444444]
ALLEGRO_UPPER = [0.558488888889, 1.727825, 1.727825, 1.727825] * 3 + [1.57075, 1.15188333333, 1.727825, 1.76273055556]


def _unit_quat(gen, *shape):
    q = torch.randn(*shape, 4, generator=gen)
    return q / q.norm(dim=-1, keepdim=True).clamp_min(1e-6)


def make_state(n: int, seed: int, c: HandConsts, fresh: bool = False, actors: int = 3) -> Dict[str, torch.Tensor]:
    gen = torch.Generator().manual_seed(seed)
    D, A, T, B = c.n_dofs, c.n_actions, c.n_tips, c.n_bodies
    U = lambda *s: torch.rand(*s, generator=gen)
    G = lambda *s: torch.randn(*s, generator=gen)
    st: Dict[str, torch.Tensor] = {}
    ar = torch.arange(n, dtype=torch.long) * actors
    st["hand_indices"], st["object_indices"], st["goal_object_indices"] = ar + 0, ar + 1, ar + 2
    lo, hi = (SHADOW_LOWER, SHADOW_UPPER) if c.kind == "shadow" else (ALLEGRO_LOWER, ALLEGRO_UPPER)
    st["dof_lower"] = torch.tensor(lo, dtype=torch.float32)
    st["dof_upper"] = torch.tensor(hi, dtype=torch.float32)
    st["dof_default"] = torch.zeros(D)
    init = torch.zeros(n, 13)
    init[:, 0:3] = torch.tensor([0.0, -0.39, 0.54]) + 0.01 * G(n, 3)      # hand pose + (0, -0.39, 0.04)-like offsets
    init[:, 6] = 1.0
    st["object_init_state"] = init
    goal = init.clone()
    goal[:, 2] -= 0.04
    st["goal_init_state"] = goal.clone()
    goal = goal.clone()
    goal[:, 3:7] = _unit_quat(gen, n)
    st["goal_states"] = goal

    root = torch.zeros(n, actors, 13)
    root[..., 0:3] = init[:, None, 0:3] + 0.05 * G(n, actors, 3)
    root[..., 3:7] = _unit_quat(gen, n, actors)
    root[..., 7:10] = 0.5 * G(n, actors, 3)
    root[..., 10:13] = 2.0 * G(n, actors, 3)
    fallen = U(n) < 0.05                                 # goal_dist >= fall_dist
    root[fallen, 1, 2] -= 0.4
    hit = U(n) < 0.1                                     # object (almost) at the goal orientation -> success branch
    root[hit, 1, 3:7] = goal[hit, 3:7] + 1e-3 * G(int(hit.sum()), 4)
    root[:, 1, 3:7] = root[:, 1, 3:7] / root[:, 1, 3:7].norm(dim=-1, keepdim=True)
    flip = U(n) < 0.1                                    # antipodal quaternions
    root[flip, 1, 3:7] = -root[flip, 1, 3:7]
    rb = torch.zeros(n, B, 13)
    rb[..., 0:3] = init[:, None, 0:3] + 0.1 * G(n, B, 3)
    rb[..., 3:7] = _unit_quat(gen, n, B)
    rb[..., 7:10] = 0.5 * G(n, B, 3)
    rb[..., 10:13] = 2.0 * G(n, B, 3)
    rb[:, c.object_body] = root[:, 1]
    st["root_state"] = root.view(-1, 13).contiguous()
    st["rigid_body_state"] = rb.view(-1, 13).contiguous()
    dof = torch.zeros(n, D, 2)
    dof[..., 0] = st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)
    dof[..., 1] = 3.0 * G(n, D)
    dof[U(n) < 0.02, :, 1] *= 20.0                       # large velocities: exercises the obs clip
    st["dof_state"] = dof.contiguous()
    st["dof_force"] = (0.3 * G(n, D)).contiguous()
    st["sensors"] = (0.2 * G(n, 6 * max(T, 1))).contiguous() if T > 0 else torch.zeros(n, 0)

    st["obs_buf"] = torch.zeros(n, c.obs_dim)
    st["states_buf"] = torch.zeros(n, c.states_dim)
    st["rew_buf"] = torch.zeros(n)
    st["timeout_buf"] = torch.zeros(n, dtype=torch.bool)
    st["actions"] = torch.zeros(n, A)
    st["rb_forces"] = torch.zeros(n, B, 3)
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * U(n) + torch.log(pr[1]))
    st["consecutive_successes"] = torch.zeros(1)
    if fresh:
        st["reset_buf"] = torch.ones(n, dtype=torch.long)
        st["reset_goal_buf"] = torch.ones(n, dtype=torch.long)
        st["progress_buf"] = torch.zeros(n, dtype=torch.long)
        st["randomize_buf"] = torch.zeros(n, dtype=torch.long)
        st["successes"] = torch.zeros(n)
        st["prev_targets"] = torch.zeros(n, D)
        st["cur_targets"] = torch.zeros(n, D)
        return st
    st["reset_buf"] = (U(n) < 0.06).long()
    st["reset_goal_buf"] = (U(n) < 0.06).long()
    st["progress_buf"] = torch.randint(0, c.max_episode_length + 2, (n,), generator=gen)
    st["randomize_buf"] = torch.randint(0, 1000, (n,), generator=gen)
    succ = torch.randint(0, 6, (n,), generator=gen).float()
    if c.max_consecutive_successes > 0:
        succ[U(n) < 0.05] = float(c.max_consecutive_successes - 1)
    st["successes"] = succ
    st["consecutive_successes"] = torch.tensor([1.7])
    st["prev_targets"] = (st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)).contiguous()
    st["cur_targets"] = st["prev_targets"].clone()
    frc = torch.zeros(n, B, 3)
    frc[:, c.object_body] = 0.1 * G(n, 3) * (U(n, 1) < 0.3)
    st["rb_forces"] = frc
    return st


def make_actions(n: int, seed: int, c: HandConsts) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed + 7919)
    return (2.0 * torch.rand(n, c.n_actions, generator=gen) - 1.0).contiguous()


def make_rnd(n: int, seed: int, c: HandConsts) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed + 104729)
    rnd = torch.rand(n, rnd_stride(c.n_dofs), generator=gen)
    rnd[:, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] = torch.randn(n, 3, generator=gen)
    return rnd.contiguous()
