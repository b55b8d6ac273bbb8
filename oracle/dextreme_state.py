"""ORACLE INFRASTRUCTURE (tests / bench CPU baseline only) -- seeded AllegroHandDextremeADR test states.

Same role as ``hand_state.py``: a dict of CPU tensors (simulator tensors in Isaac Gym's layout, every per-env task tensor, and the
per-env ADR / affine parameters that ``apply_randomizations`` would have sampled) that can be injected into the reference class,
fed to the restatement (``dextreme_oracle``) and loaded into the CUDA task (``tests/test_dextreme_parity_gpu.py``).
"""
from __future__ import annotations

from typing import Dict

import torch

from .dextreme_oracle import (ADR_KEYS, ADR_N_PARAMS, ADR_TENSOR_KEYS, ARND_AFFINE, ARND_STRIDE, OBS_DIM, PRND_CUBE_MASK, PRND_DELAY, PRND_DOF_WHITE, PRND_POSE_WHITE, PRND_STRIDE, RND_ACTION_WHITE,
                              RND_FORCE_NORMAL, RND_STRIDE, DexConsts)
from .hand_state import ALLEGRO_LOWER, ALLEGRO_UPPER, _unit_quat


def make_state(n: int, seed: int, c: DexConsts, fresh: int = 0, actors: int = 3) -> Dict[str, torch.Tensor]:
    """``fresh``: 0 = mid-episode mix, 1 = just constructed (reset_buf ones, nothing applied yet), 2 = every env resets now."""
    gen = torch.Generator().manual_seed(seed)
    D, T, B, L = c.n_dofs, c.n_tips, c.n_bodies, c.latency_max + 1
    U = lambda *s: torch.rand(*s, generator=gen)
    G = lambda *s: torch.randn(*s, generator=gen)
    st: Dict[str, torch.Tensor] = {}
    ar = torch.arange(n, dtype=torch.long) * actors
    st["hand_indices"], st["object_indices"], st["goal_object_indices"] = ar + 0, ar + 1, ar + 2
    st["dof_lower"] = torch.tensor(ALLEGRO_LOWER, dtype=torch.float32)
    st["dof_upper"] = torch.tensor(ALLEGRO_UPPER, dtype=torch.float32)
    st["dof_default"] = torch.zeros(D)
    init = torch.zeros(n, 13)
    init[:, 0:3] = torch.tensor([0.0, -0.15, 0.56]) + 0.01 * G(n, 3)      # hand start + (0, startObjectPoseDY, startObjectPoseDZ)
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
    dof[U(n) < 0.02, :, 1] *= 30.0                       # large velocities: exercises the +-50 clip
    st["dof_state"] = dof.contiguous()
    st["dof_force"] = (0.3 * G(n, D)).contiguous()
    st["sensors"] = (0.2 * G(n, 6 * T)).contiguous()

    # per-env ADR tensors / affine parameters (what apply_randomizations samples, allegro_hand_dextreme.py:1392-1416)
    st["action_latency"] = torch.randint(0, c.latency_max + 1, (n,), generator=gen)
    st["cube_pose_refresh_rate"] = torch.randint(1, 7, (n,), generator=gen)
    st["cube_pose_refresh_offset"] = (U(n) * st["cube_pose_refresh_rate"].float() - 0.5).round().long()
    for k in ADR_KEYS:
        v = 0.3 * U(n) if k.endswith("prob") else 0.25 * U(n)
        v[U(n) < 0.2] = 0.0                              # the v > 0 gate of sample_gaussian_adr
        st["adr_" + k] = v
    st["adr_rna_alpha"] = 0.5 * U(n)
    for k in ADR_TENSOR_KEYS:                            # the remaining tensorised ADR parameters (only resampling reads them)
        if "adr_" + k not in st:
            st["adr_" + k] = 0.2 * U(n)
    st["worker_types"] = (U(n) < 0.6).long()
    st["adr_modes"] = torch.randint(0, 2 * ADR_N_PARAMS, (n,), generator=gen)
    st["affine_actions_scaling"] = 1.0 + 0.05 * G(n, D)
    st["affine_actions_additive"] = 0.05 * G(n, D)
    st["affine_cube_pose_scaling"] = 1.0 + 0.05 * G(n, 7)
    st["affine_cube_pose_additive"] = 0.05 * G(n, 7)
    st["affine_dof_pos_scaling"] = 1.0 + 0.05 * G(n, D)
    st["affine_dof_pos_additive"] = 0.05 * G(n, D)
    st["cube_random_params"] = torch.stack([0.95 + 0.1 * U(n), 0.8 + 0.4 * U(n), 0.4 + 0.4 * U(n)], dim=1)
    st["hand_random_params"] = 0.95 + 0.1 * U(n, 1)
    st["gravity_vec"] = torch.tensor([0.0, 0.0, -9.81]) + 0.3 * G(n, 3)

    st["rew_buf"] = torch.zeros(n)
    st["timeout_buf"] = torch.zeros(n, dtype=torch.bool)
    st["actions"] = torch.zeros(n, D)
    st["rb_forces"] = torch.zeros(n, B, 3)
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * U(n) + torch.log(pr[1]))
    st["consecutive_successes"] = torch.zeros(1)
    st["random_cube_poses"] = torch.zeros(n, 7)
    st["object_pose"] = st["root_state"][st["object_indices"], 0:7].clone()
    st["object_pose_wrt_wrist"] = torch.zeros(n, 7)
    st["object_pose_wrt_wrist"][:, 6] = 1.0
    st["obs_object_pose"] = st["object_pose"].clone()
    st["obs_object_pose_freq"] = st["object_pose"].clone()
    st["dof_delta"] = torch.zeros(n, D)
    st["episode_cumulative"] = torch.zeros(8, n)
    st["obs_rows"] = torch.zeros(n, OBS_DIM)
    st["frame"] = torch.tensor(0, dtype=torch.long)
    if fresh:
        st["reset_buf"] = torch.ones(n, dtype=torch.long)
        st["reset_goal_buf"] = torch.ones(n, dtype=torch.long)
        st["apply_reset_buf"] = torch.full((n,), 1 if fresh == 2 else 0, dtype=torch.long)
        st["progress_buf"] = torch.zeros(n, dtype=torch.long)
        st["hold_count_buf"] = torch.zeros(n, dtype=torch.long)
        st["successes"] = torch.zeros(n)
        st["best_rotation_dist"] = -torch.ones(n)
        for k in ("prev_targets", "cur_targets", "prev_prev_targets", "prev_actions"):
            st[k] = torch.zeros(n, D)
        st["prev_actions_queue"] = torch.zeros(n, L, D)
        return st
    st["frame"] = torch.tensor(1234 + seed % 7, dtype=torch.long)
    st["reset_buf"] = (U(n) < 0.1).long()
    st["apply_reset_buf"] = st["reset_buf"] * (U(n) < 0.5).long()          # half of the pending resets were randomised a step ago
    st["reset_goal_buf"] = (U(n) < 0.06).long()
    st["progress_buf"] = torch.randint(0, c.max_episode_length + 2, (n,), generator=gen)
    st["progress_buf"][U(n) < 0.05] = 0                                    # refreshes the action queue
    st["hold_count_buf"] = torch.randint(0, 3, (n,), generator=gen)
    succ = torch.randint(0, 6, (n,), generator=gen).float()
    succ[U(n) < 0.05] = float(c.max_consecutive_successes - 1)
    st["successes"] = succ
    st["consecutive_successes"] = torch.tensor([1.7])
    best = 3.0 * U(n)
    best[U(n) < 0.2] = -1.0
    st["best_rotation_dist"] = best
    span = st["dof_upper"] - st["dof_lower"]
    st["prev_targets"] = (st["dof_lower"] + span * U(n, D)).contiguous()
    st["cur_targets"] = (st["prev_targets"] + 0.05 * G(n, D)).contiguous()
    st["prev_prev_targets"] = (st["dof_lower"] + span * U(n, D)).contiguous()
    st["prev_actions"] = 2.0 * U(n, D) - 1.0
    st["prev_actions_queue"] = 2.0 * U(n, L, D) - 1.0
    st["object_pose"] = st["object_pose"] + 0.01 * G(n, 7)                # the gather of the previous compute_observations
    st["obs_object_pose"] = torch.cat([init[:, 0:3] + 0.05 * G(n, 3), _unit_quat(gen, n)], dim=-1)
    st["obs_object_pose_freq"] = torch.cat([init[:, 0:3] + 0.05 * G(n, 3), _unit_quat(gen, n)], dim=-1)
    frc = torch.zeros(n, B, 3)
    frc[:, c.object_body] = 0.1 * G(n, 3) * (U(n, 1) < 0.3)
    st["rb_forces"] = frc
    return st


def make_actions(n: int, seed: int, c: DexConsts) -> torch.Tensor:
    gen = torch.Generator().manual_seed(seed + 7919)
    return (2.0 * torch.rand(n, c.n_dofs, generator=gen) - 1.0).contiguous()


def make_rnd(n: int, seed: int, c: DexConsts) -> torch.Tensor:
    """Pre-step table [n, RND_STRIDE]: uniforms, with N(0,1) in the white-noise and force slots."""
    gen = torch.Generator().manual_seed(seed + 104729)
    rnd = torch.rand(n, RND_STRIDE, generator=gen)
    rnd[:, RND_ACTION_WHITE:RND_ACTION_WHITE + c.n_dofs] = torch.randn(n, c.n_dofs, generator=gen)
    rnd[:, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] = torch.randn(n, 3, generator=gen)
    return rnd.contiguous()


def make_prnd(n: int, seed: int, c: DexConsts) -> torch.Tensor:
    """Post-step table [n, PRND_STRIDE]."""
    gen = torch.Generator().manual_seed(seed + 1299709)
    rnd = torch.rand(n, PRND_STRIDE, generator=gen)
    rnd[:, PRND_POSE_WHITE:PRND_POSE_WHITE + 7] = torch.randn(n, 7, generator=gen)
    rnd[:, PRND_DOF_WHITE:PRND_DOF_WHITE + c.n_dofs] = torch.randn(n, c.n_dofs, generator=gen)
    return rnd.contiguous()


def make_arnd(n: int, seed: int, c: DexConsts) -> torch.Tensor:
    """ADR-resampling table [n, ARND_STRIDE]: uniforms, N(0,1) in the affine slots."""
    gen = torch.Generator().manual_seed(seed + 15485863)
    rnd = torch.rand(n, ARND_STRIDE, generator=gen)
    rnd[:, ARND_AFFINE:] = torch.randn(n, ARND_STRIDE - ARND_AFFINE, generator=gen)
    return rnd.contiguous()
