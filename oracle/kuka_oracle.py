"""ORACLE (test infrastructure, not product code) -- CPU restatement of the AllegroKuka hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline leg may import this.

A functional, torch-on-CPU restatement of what the reference does per step for the single-arm
AllegroKuka tasks, written op-for-op in the reference's association order so that fp32 results
are bit-comparable with the reference's eager code:

* ``pre_step``   <- ``allegro_kuka_base.py:1355-1424`` (pre_physics_step), ``:1246-1353`` (reset_idx),
                    ``:1191-1224`` (reset_target_pose / reset_object_pose), ``:1178-1189`` (get_random_quat),
                    ``allegro_kuka_reorientation.py:104-128`` (_reset_target)
* ``post_step``  <- ``allegro_kuka_base.py:1426-1443`` (post_physics_step), ``:991-1172`` (observations),
                    ``:759-930`` (rewards, resets), ``allegro_kuka_reorientation.py:97-102``,
                    ``allegro_kuka_utils.py:128-158`` (true objective), ``vec_task.py:424`` (timeout_buf)
* quaternion helpers <- ``utils/torch_jit_utils.py:80-90`` (quat_rotate), ``:233-240`` (scale/unscale)

Parity pinning: the reference has no golden vectors for this path.  This file is
pinned against the reference itself: ``tests/test_oracle_vs_reference.py`` runs the reference's real
classes (through ``oracle/reference_loader.py``) and this restatement on the same inputs whenever
``/root/reference`` is present, and ``tests/golden/kuka_*.npz`` hold reference-generated vectors
(made by ``oracle/make_golden.py``) that travel to machines without the reference.

Random numbers: instead of torch's global generator the functions consume ``rnd``, a per-env table
of raw draws laid out as ``HA_KUKA_RND_*`` in ``include/handarm_b200.h`` (the reference loader feeds the
same numbers to the reference by patching ``torch.rand`` / ``torch.randn``).
"""
from __future__ import annotations

import dataclasses
import math
from typing import Dict, List, Tuple

import numpy as np
import torch

Tensor = torch.Tensor

# slot layout of the per-env random table (include/handarm_b200.h, HA_KUKA_RND_*)
RND_GOAL_POS, RND_GOAL_ROT, RND_RESET_GOAL_POS, RND_RESET_GOAL_ROT = 0, 3, 6, 9
RND_OBJ_POS, RND_OBJ_ROT, RND_FORCE_PROB, RND_FORCE_TEST, RND_FORCE_NORMAL, RND_SIDE, RND_DOF_POS = 12, 15, 18, 19, 20, 23, 25


def rnd_stride(n_dofs: int) -> int:
    return 25 + 2 * n_dofs


EPISODE_TERMS = ["raw_fingertip_delta_rew", "raw_hand_delta_penalty", "raw_lifting_rew", "raw_keypoint_rew", "fingertip_delta_rew",
                 "hand_delta_penalty", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "bonus_rew", "kuka_actions_penalty",
                 "allegro_actions_penalty"]
STEP_TERMS = ["fingertip_delta_rew", "hand_delta_penalty", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "kuka_actions_penalty",
              "allegro_actions_penalty", "bonus_rew"]


@dataclasses.dataclass
class KukaConsts:
    """Host scalars of the task (python floats, exactly the reference's attributes)."""
    n_arm_dofs: int = 7
    n_dofs: int = 23
    n_tips: int = 4
    palm_body: int = 7
    tip_bodies: Tuple[int, ...] = (11, 15, 19, 23)
    object_body: int = 24
    palm_offset: Tuple[float, ...] = (-0.0, -0.02, 0.16)
    tip_offsets: Tuple[Tuple[float, ...], ...] = ((0.05, 0.005, 0), (0.05, 0.005, 0), (0.05, 0.005, 0), (0.06, 0.005, 0))
    dt: float = 0.01667
    clamp_abs_observations: float = 10.0
    max_episode_length: int = 600
    max_consecutive_successes: int = 50
    success_steps: int = 1
    success_tolerance: float = 0.075
    initial_tolerance: float = 0.075
    target_tolerance: float = 0.01
    keypoint_scale: float = 1.5
    lifting_bonus_threshold: float = 0.15
    lifting_bonus: float = 300.0
    lifting_rew_scale: float = 20.0
    keypoint_rew_scale: float = 200.0
    distance_delta_rew_scale: float = 50.0
    reach_goal_bonus: float = 1000.0
    kuka_actions_penalty_scale: float = 0.003
    allegro_actions_penalty_scale: float = 0.0003
    hand_dof_speed_scale: float = 10.0
    act_moving_average: float = 1.0
    force_scale: float = 2.0
    force_decay: float = 0.99
    force_decay_interval: float = 0.08
    force_prob_range: Tuple[float, float] = (0.001, 0.1)
    object_mass: float = 0.05
    reset_position_noise: Tuple[float, float, float] = (0.1, 0.1, 0.02)
    reset_dof_pos_noise_arm: float = 0.1
    reset_dof_pos_noise_fingers: float = 0.1
    reset_dof_vel_noise: float = 0.5
    target_volume_origin: Tuple[float, ...] = (0, 0.05, 0.8)
    target_volume_extent: Tuple[Tuple[float, float], ...] = ((-0.4, 0.4), (-0.05, 0.3), (-0.12, 0.25))
    extra_far_reset: bool = True          # reorientation's _extra_reset_rules


# ---------------------------------------------------------------------------------------------
# math helpers (torch_jit_utils.py)
# ---------------------------------------------------------------------------------------------
def quat_rotate(q: Tensor, v: Tensor) -> Tensor:
    """torch_jit_utils.py:80-90; q [..,4] xyzw, v [..,3]."""
    qw = q[..., 3:4]
    qv = q[..., 0:3]
    a = v * (2.0 * qw ** 2 - 1.0)
    b = torch.cross(qv, v, dim=-1) * qw * 2.0
    dot = (qv[..., 0] * v[..., 0] + qv[..., 1] * v[..., 1] + qv[..., 2] * v[..., 2]).unsqueeze(-1)
    c = qv * dot * 2.0
    return a + b + c


def unscale(x, lo, hi):
    return (2.0 * x - hi - lo) / (hi - lo)


def scale(x, lo, hi):
    return 0.5 * (x + 1.0) * (hi - lo) + lo


def random_quat(uvw: Tensor) -> Tensor:
    """allegro_kuka_base.py:1178-1189."""
    two_pi = 2 * np.pi
    s1, s2 = torch.sqrt(1.0 - uvw[:, 0]), torch.sqrt(uvw[:, 0])
    qw = s1 * torch.sin(two_pi * uvw[:, 1])
    qx = s1 * torch.cos(two_pi * uvw[:, 1])
    qy = s2 * torch.sin(two_pi * uvw[:, 2])
    qz = s2 * torch.cos(two_pi * uvw[:, 2])
    return torch.stack((qx, qy, qz, qw), dim=-1)


def true_objective_coeffs(tol: float, initial: float, target: float) -> Tuple[float, float]:
    """allegro_kuka_utils.py:118-158."""
    offset = (initial - tol) / (initial - target) if initial > target else 1.0
    return (0.01 if tol > target else 1.0), offset
