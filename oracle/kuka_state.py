"""ORACLE INFRASTRUCTURE (tests / bench CPU baseline only) -- seeded AllegroKuka test states.

``make_state`` builds a complete mid-episode state as a dict of CPU tensors (simulator tensors in
Isaac Gym's layout + every per-env task tensor) that can be (a) injected into the reference classes
(``reference_loader.make_reference_kuka``), (b) fed to the restatement (``kuka_oracle``) and (c) loaded
into the CUDA task (``tests/kuka_harness.py``).  Distributions follow SURVEY.md 8d, with the branch
probabilities raised so that small N still exercises every branch.
"""
from __future__ import annotations

from typing import Dict

import numpy as np
import torch

from .kuka_oracle import KukaConsts, rnd_stride

KUKA_LOWER = [-2.96705972839, -2.09439510239, -2.96705972839, -2.09439510239, -2.96705972839, -2.09439510239, -3.05432619099] + [
    -0.558488888889, -0.279244444444, -0.279244444444, -0.279244444444] * 3 + [0.279244444444, -0.331602777778, -0.279244444444,
                                                                              -0.279244444444]
KUKA_UPPER = [2.96705972839, 2.09439510239, 2.96705972839, 2.09439510239, 2.96705972839, 2.09439510239, 3.05432619099] + [
    0.558488888889, 1.727825, 1.727825, 1.727825] * 3 + [1.57075, 1.15188333333, 1.727825, 1.76273055556]


def _unit_quat(gen, *shape):
    q = torch.randn(*shape, 4, generator=gen)
    return q / q.norm(dim=-1, keepdim=True).clamp_min(1e-6)


def make_state(n: int, seed: int = 0, c: KukaConsts = KukaConsts(), n_keypoints: int = 4, actors: int = 4, bodies: int = 27,
               fresh: bool = False) -> Dict[str, torch.Tensor]:
    """``fresh=True`` gives the just-constructed state (reset_buf all ones); otherwise a random mid-episode one."""
    gen = torch.Generator().manual_seed(seed)
    single_keypoint = getattr(c, "subtask", "reorientation") != "reorientation"      # regrasping / throw: one keypoint at the object centre
    if single_keypoint:
        n_keypoints = 1
    D, T, K = c.n_dofs, c.n_tips, n_keypoints
    obs_dim = 2 * D + 3 + 10 + 10 + 3 * T + 6 * K + 3 + 1 + T + 4
    U = lambda *s: torch.rand(*s, generator=gen)
    G = lambda *s: torch.randn(*s, generator=gen)
    st: Dict[str, torch.Tensor] = {}
    ar = torch.arange(n, dtype=torch.long) * actors
    st["allegro_hand_indices"], st["object_indices"], st["goal_object_indices"] = ar + 0, ar + 1, ar + 3
    st["dof_lower"] = torch.tensor(KUKA_LOWER, dtype=torch.float32)
    st["dof_upper"] = torch.tensor(KUKA_UPPER, dtype=torch.float32)
    dflt = torch.zeros(D)
    dflt[:7] = torch.tensor([-1.571, 1.571, -0.000, 1.376, -0.000, 1.485, 2.358])
    st["dof_default"] = dflt
    init = torch.zeros(n, 13)
    init[:, 0:3] = torch.tensor([0.0, 0.0, 0.63])
    init[:, 6] = 1.0
    st["object_init_state"] = init
    goal = init.clone()
    goal[:, 0:3] = torch.tensor([0.0, 0.05, 0.8]) + 0.15 * G(n, 3)
    goal[:, 3:7] = _unit_quat(gen, n)
    st["goal_states"] = goal
    scales = torch.tensor([[1.0, 1.0, 1.0], [4.0, 0.75, 1.0], [1.0, 2.5, 1.0], [1.25, 0.66, 3.0], [0.5, 2.0, 1.5]])
    st["object_scales"] = scales[torch.arange(n) % len(scales)].contiguous()
    corners = torch.tensor([[1, 1, 1], [1, 1, -1], [-1, -1, 1], [-1, -1, -1]], dtype=torch.float32)[:K]
    if single_keypoint:
        corners = torch.zeros(1, 3)
    st["object_keypoint_offsets"] = (corners[None] * (st["object_scales"][:, None, :] * 0.05 * c.keypoint_scale / 2)).contiguous()
    st["finger_rew_coeffs"] = torch.ones(n, T)

    # simulator tensors
    root = torch.zeros(n, actors, 13)
    root[..., 0:3] = torch.tensor([0.0, 0.0, 0.63]) + 0.1 * G(n, actors, 3)
    root[..., 3:7] = _unit_quat(gen, n, actors)
    root[..., 7:10] = 0.5 * G(n, actors, 3)
    root[..., 10:13] = 2.0 * G(n, actors, 3)
    root[U(n) < 0.04, 1, 2] = 0.05                      # fallen objects
    hit = U(n) < 0.08                                   # objects (almost) at the goal pose
    root[hit, 1, 0:7] = goal[hit, 0:7] + 1e-3 * G(int(hit.sum()), 7)
    rb = torch.zeros(n, bodies, 13)
    rb[..., 0:3] = root[:, 1:2, 0:3] + 0.12 * G(n, bodies, 3)
    rb[..., 3:7] = _unit_quat(gen, n, bodies)
    rb[..., 7:10] = 0.5 * G(n, bodies, 3)
    rb[..., 10:13] = 2.0 * G(n, bodies, 3)
    rb[U(n) < 0.02, :, 0] += 2.0                        # hands far from the object
    rb[:, c.object_body] = root[:, 1]
    st["root_state"] = root.view(-1, 13).contiguous()
    st["rigid_body_state"] = rb.view(-1, 13).contiguous()
    dof = torch.zeros(n, D, 2)
    dof[..., 0] = st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)
    dof[..., 1] = G(n, D)
    st["dof_state"] = dof.contiguous()

    # task state
    st["obs_buf"] = torch.zeros(n, obs_dim)
    st["rew_buf"] = torch.zeros(n)
    st["timeout_buf"] = torch.zeros(n, dtype=torch.bool)
    st["randomize_buf"] = torch.zeros(n, dtype=torch.long)
    st["prev_targets"] = torch.zeros(n, D)
    st["cur_targets"] = torch.zeros(n, D)
    st["actions"] = torch.zeros(n, D)
    st["rb_forces"] = torch.zeros(n, bodies, 3)
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * U(n) + torch.log(pr[1]))
    st["rewards_episode"] = torch.zeros(12, n)
    st["episode_cumulative"] = torch.zeros(8, n)
    if fresh:
        st["reset_buf"] = torch.ones(n, dtype=torch.long)
        st["reset_goal_buf"] = torch.ones(n, dtype=torch.long)
        st["progress_buf"] = torch.zeros(n, dtype=torch.long)
        st["successes"] = torch.zeros(n)
        st["prev_episode_successes"] = torch.zeros(n)
        st["true_objective"] = torch.zeros(n)
        st["prev_episode_true_objective"] = torch.zeros(n)
        st["near_goal_steps"] = torch.zeros(n, dtype=torch.int32)
        st["lifted_object"] = torch.zeros(n, dtype=torch.bool)
        st["closest_keypoint_max_dist"] = -torch.ones(n)
        st["closest_fingertip_dist"] = -torch.ones(n, T)
        st["furthest_hand_dist"] = -torch.ones(n)
        return st
    st["reset_buf"] = (U(n) < 0.05).long()
    st["reset_goal_buf"] = (U(n) < 0.05).long()
    st["progress_buf"] = torch.randint(0, c.max_episode_length + 2, (n,), generator=gen)
    st["randomize_buf"] = torch.randint(0, 1000, (n,), generator=gen)
    succ = torch.randint(0, 8, (n,), generator=gen).float()
    succ[U(n) < 0.03] = float(c.max_consecutive_successes - 1)
    st["successes"] = succ
    st["prev_episode_successes"] = torch.randint(0, 8, (n,), generator=gen).float()
    st["true_objective"] = succ * 0.01
    st["prev_episode_true_objective"] = st["prev_episode_successes"] * 0.01
    st["near_goal_steps"] = torch.randint(0, max(c.success_steps, 1), (n,), generator=gen).to(torch.int32)
    st["lifted_object"] = U(n) < 0.4
    ckp = 0.05 + 0.3 * U(n)
    ckp[U(n) < 0.2] = -1.0
    st["closest_keypoint_max_dist"] = ckp
    ctip = 0.05 + 0.3 * U(n, T)
    ctip[U(n) < 0.2] = -1.0
    st["closest_fingertip_dist"] = ctip
    fur = 0.1 + 0.3 * U(n)
    fur[ctip[:, 0] < 0] = -1.0
    st["furthest_hand_dist"] = fur
    st["prev_targets"] = (st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)).contiguous()
    st["cur_targets"] = st["prev_targets"].clone()
    frc = torch.zeros(n, bodies, 3)
    frc[:, c.object_body] = 0.1 * G(n, 3) * (U(n, 1) < 0.3)
    st["rb_forces"] = frc
    st["rewards_episode"] = G(12, n).contiguous()
    return st


def make_actions(n: int, seed: int, c: KukaConsts = KukaConsts()) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed + 7919)
    return (2.0 * torch.rand(n, c.n_dofs, generator=gen) - 1.0).contiguous()


def make_rnd(n: int, seed: int, c: KukaConsts = KukaConsts()) -> torch.Tensor:
    """Per-env table of raw draws: uniforms everywhere, N(0,1) in the force-normal slots."""
    from .kuka_oracle import RND_FORCE_NORMAL
    gen = torch.Generator().manual_seed(seed + 104729)
    rnd = torch.rand(n, rnd_stride(c.n_dofs), generator=gen)
    rnd[:, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] = torch.randn(n, 3, generator=gen)
    return rnd.contiguous()


def clone_state(st: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    return {k: v.clone() for k, v in st.items()}
