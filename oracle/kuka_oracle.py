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
    subtask: str = "reorientation"        # "reorientation" | "regrasping" | "throw" (env.subtask; the task classes of the same names)


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


# ---------------------------------------------------------------------------------------------
# pre-physics step
# ---------------------------------------------------------------------------------------------
def _reset_object_pose(st: Dict[str, Tensor], c: KukaConsts, ids: Tensor, rnd: Tensor) -> None:
    """allegro_kuka_base.py:1198-1224 for env ids ``ids`` (draws: RND_OBJ_POS, RND_OBJ_ROT)."""
    rows = st["object_indices"][ids]
    noise = 2.0 * rnd[ids, RND_OBJ_POS:RND_OBJ_POS + 3] + (-1.0)
    new_row = st["object_init_state"][ids].clone()
    for k in range(3):
        new_row[:, k] = st["object_init_state"][ids, k] + c.reset_position_noise[k] * noise[:, k]
    new_row[:, 3:7] = random_quat(rnd[ids, RND_OBJ_ROT:RND_OBJ_ROT + 3])
    new_row[:, 7:13] = 0.0
    st["root_state"][rows] = new_row
    st["closest_fingertip_dist"][ids] = -1.0
    st["furthest_hand_dist"][ids] = -1.0


def _reset_target(st: Dict[str, Tensor], c: KukaConsts, ids: Tensor, rnd: Tensor, slot: int) -> None:
    """reset_target_pose (allegro_kuka_base.py:1191-1196) with the subtask's ``_reset_target`` for env ids ``ids``; ``slot`` is the
    first of the (up to six) raw draws of the goal: RND_GOAL_POS for a goal-only reset, RND_RESET_GOAL_POS inside reset_idx.
      reorientation  allegro_kuka_reorientation.py:104-128: position in the target volume + random rotation
      regrasping     allegro_kuka_regrasping.py:76-100: position only; the object goes back to the table, lifted_object = False
      throw          allegro_kuka_throw.py:90-117: bucket left / right of the table; goal = bucket + 5 cm; object back to the table"""
    rows = st["goal_object_indices"][ids]
    if c.subtask == "throw":
        u = rnd[ids, slot:slot + 4]
        left_right = 2.0 * u[:, 0:1] + (-1.0)
        x_pos = torch.where(left_right > 0, 0.5 * torch.ones_like(left_right), -0.5 * torch.ones_like(left_right))
        x_pos = x_pos + torch.sign(left_right) * (0.4 * u[:, 1:2] + 0.0)
        y_pos = (0.7 - (-1.0)) * u[:, 2:3] + (-1.0)
        z_pos = (1.0 - 0.0) * u[:, 3:4] + 0.0
        st["root_state"][rows, 0:1] = x_pos
        st["root_state"][rows, 1:2] = y_pos
        st["root_state"][rows, 2:3] = z_pos
        st["goal_states"][ids, 0:1] = x_pos
        st["goal_states"][ids, 1:2] = y_pos
        st["goal_states"][ids, 2:3] = z_pos + 0.05
    else:
        org = torch.tensor(c.target_volume_origin, dtype=torch.float32)
        ext = torch.tensor(c.target_volume_extent, dtype=torch.float32)
        lo = org + ext[:, 0]
        size = (org + ext[:, 1]) - lo
        pos = lo + rnd[ids, slot:slot + 3] * size
        if c.subtask == "reorientation":
            pose = torch.cat((pos, random_quat(rnd[ids, slot + 3:slot + 6])), dim=-1)
            st["goal_states"][ids, 0:7] = pose
            st["root_state"][rows, 0:7] = pose
            st["root_state"][rows, 7:13] = 0.0
        else:
            st["goal_states"][ids, 0:3] = pos
            st["root_state"][rows, 0:3] = pos
    if c.subtask != "reorientation":
        _reset_object_pose(st, c, ids, rnd)
        st["lifted_object"][ids] = False
    st["reset_goal_buf"][ids] = 0
    st["near_goal_steps"][ids] = 0
    st["closest_keypoint_max_dist"][ids] = -1.0


def load_initial_states(st: Dict[str, Tensor], ids: Tensor, initial: Dict[str, object]) -> None:
    """The ``should_load_initial_states`` branch of ``reset_idx`` (allegro_kuka_base.py:1292-1313): the reset envs, in ascending
    order, take consecutive recorded states -- every DOF state and the object's root row; the targets keep the re-randomised
    position.  ``initial`` = {"root": [T, actors, 13], "dof": [T, dofs_per_env, 2], "idx": cursor (mutated)}."""
    T, k = initial["root"].shape[0], len(ids)
    if k > T:
        return                                        # "Not enough initial states to load"
    if initial["idx"] + k > T:
        initial["idx"] = 0
    i0 = initial["idx"]
    st["dof_state"][ids] = initial["dof"][i0:i0 + k].clone()
    N = st["reset_buf"].shape[0]
    slot = int(st["object_indices"][0])               # cube_object_idx = self.object_indices[0] (:1308)
    st["root_state"].view(N, -1, 13)[ids, slot] = initial["root"][i0:i0 + k, slot].clone()
    initial["idx"] = i0 + k


def pre_step(st: Dict[str, Tensor], c: KukaConsts, actions: Tensor, rnd: Tensor, initial: Dict[str, object] = None) -> Dict[str, Tensor]:
    """One ``pre_physics_step``; mutates ``st`` in place and returns the simulator-facing index lists."""
    D, A = c.n_dofs, c.n_arm_dofs
    lo, hi = st["dof_lower"], st["dof_upper"]
    st["actions"] = actions.clone()
    reset_ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    root_lists: List[Tensor] = []

    # goal-only resets (always executed, also for an empty id list)
    _reset_target(st, c, goal_ids, rnd, RND_GOAL_POS)
    root_lists.append(st["goal_object_indices"][goal_ids])
    if c.subtask != "reorientation":
        root_lists.append(st["object_indices"][goal_ids])

    hand_rows = torch.zeros(0, dtype=torch.int32)
    if len(reset_ids) > 0:
        ids = reset_ids
        _reset_target(st, c, ids, rnd, RND_RESET_GOAL_POS)
        root_lists.append(st["goal_object_indices"][ids])
        st["rb_forces"][ids] = 0.0
        _reset_object_pose(st, c, ids, rnd)      # (the object draws a regrasping / throw _reset_target made just before are overwritten)
        hand_rows = st["allegro_hand_indices"][ids].to(torch.int32)
        # force probability (:1265-1269)
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        st["random_force_prob"][ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, RND_FORCE_PROB] + torch.log(pr[1]))
        # DOF state (:1271-1290)
        dflt = st["dof_default"]
        dmax, dmin = hi - dflt, lo - dflt
        coeff = torch.zeros_like(dflt)
        coeff[0:A] = c.reset_dof_pos_noise_arm
        coeff[A:D] = c.reset_dof_pos_noise_fingers
        pos = dflt + coeff * (dmin + (dmax - dmin) * rnd[ids, RND_DOF_POS:RND_DOF_POS + D])
        vel = c.reset_dof_vel_noise * (2.0 * rnd[ids, RND_DOF_POS + D:RND_DOF_POS + 2 * D] + (-1.0))
        st["dof_state"][ids, :D, 0] = pos
        st["dof_state"][ids, :D, 1] = vel
        st["prev_targets"][ids, :D] = pos
        st["cur_targets"][ids, :D] = pos
        if initial is not None:
            load_initial_states(st, ids, initial)
        root_lists.append(st["object_indices"][ids])
        # episode bookkeeping (:1323-1347)
        st["progress_buf"][ids] = 0
        st["reset_buf"][ids] = 0
        st["prev_episode_successes"][ids] = st["successes"][ids]
        st["successes"][ids] = 0.0
        st["prev_episode_true_objective"][ids] = st["true_objective"][ids]
        st["true_objective"][ids] = 0.0
        st["lifted_object"][ids] = False
        st["closest_keypoint_max_dist"][ids] = -1.0
        st["closest_fingertip_dist"][ids] = -1.0
        st["furthest_hand_dist"][ids] = -1.0
        st["near_goal_steps"][ids] = 0
        st["rewards_episode"][:, ids] = 0.0

    root_rows = torch.unique(torch.cat(root_lists).to(torch.int32))       # (:1229-1244)

    # action -> targets (:1377-1397)
    cur = st["cur_targets"]
    prev = st["prev_targets"]
    hand = scale(actions[:, A:D], lo[A:D], hi[A:D])
    hand = c.act_moving_average * hand + (1.0 - c.act_moving_average) * prev[:, A:D]
    cur[:, A:D] = torch.max(torch.min(hand, hi[A:D]), lo[A:D])
    arm = prev[:, :A] + c.hand_dof_speed_scale * c.dt * actions[:, :A]
    cur[:, :A] = torch.max(torch.min(arm, hi[:A]), lo[:A])
    prev[:, :] = cur

    # random forces on the object body (:1401-1414)
    if c.force_scale > 0.0:
        st["rb_forces"] *= torch.pow(torch.tensor(c.force_decay, dtype=torch.float32), c.dt / c.force_decay_interval)
        hit = (rnd[:, RND_FORCE_TEST] < st["random_force_prob"]).nonzero(as_tuple=False).squeeze(-1)
        masses = torch.tensor([c.object_mass], dtype=torch.float32)
        st["rb_forces"][hit, c.object_body, :] = rnd[hit, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] * masses * c.force_scale
    return {"root_rows": root_rows, "hand_rows": hand_rows}


# ---------------------------------------------------------------------------------------------
# post-physics step
# ---------------------------------------------------------------------------------------------
def _geometry(st: Dict[str, Tensor], c: KukaConsts) -> Dict[str, Tensor]:
    """Gathers + keypoint / fingertip geometry of compute_observations (allegro_kuka_base.py:1002-1079)."""
    N = st["reset_buf"].shape[0]
    obj = st["root_state"][st["object_indices"]]
    rb = st["rigid_body_state"].view(N, -1, 13)
    palm = rb[:, c.palm_body]
    tips = rb[:, list(c.tip_bodies)]
    g: Dict[str, Tensor] = {"obj": obj, "palm": palm}
    palm_off = torch.tensor(c.palm_offset, dtype=torch.float32).repeat(N, 1)
    g["palm_center"] = palm[:, 0:3] + quat_rotate(palm[:, 3:7], palm_off)
    tip_off = torch.tensor(np.asarray(c.tip_offsets, dtype=np.float32)).repeat(N, 1, 1)
    tip_pos = torch.zeros(N, c.n_tips, 3)
    for i in range(c.n_tips):
        tip_pos[:, i] = tips[:, i, 0:3] + quat_rotate(tips[:, i, 3:7], tip_off[:, i])
    g["tip_rel_obj"] = tip_pos - obj[:, None, 0:3]
    g["tip_dist"] = torch.norm(g["tip_rel_obj"], dim=-1)
    g["tip_rel_palm"] = tip_pos - g["palm_center"][:, None]
    K = st["object_keypoint_offsets"].shape[1]
    okp = torch.zeros(N, K, 3)
    gkp = torch.zeros(N, K, 3)
    goal = st["goal_states"]
    for k in range(K):
        okp[:, k] = obj[:, 0:3] + quat_rotate(obj[:, 3:7], st["object_keypoint_offsets"][:, k])
        gkp[:, k] = goal[:, 0:3] + quat_rotate(goal[:, 3:7], st["object_keypoint_offsets"][:, k])
    g["kp_rel_goal"] = okp - gkp
    g["kp_rel_palm"] = okp - g["palm_center"][:, None]
    g["kp_max"] = torch.norm(g["kp_rel_goal"], dim=-1).max(dim=-1).values
    return g


def _fill_obs(st: Dict[str, Tensor], c: KukaConsts, g: Dict[str, Tensor]) -> int:
    """compute_full_state (allegro_kuka_base.py:1091-1172); returns the reward slot offset."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    dof = st["dof_state"].view(N, -1, 2)[:, :D]
    parts = [
        unscale(dof[..., 0], st["dof_lower"], st["dof_upper"]),
        dof[..., 1],
        g["palm_center"],
        g["palm"][:, 3:13],
        g["obj"][:, 3:13],
        g["tip_rel_palm"].reshape(N, -1),
        g["kp_rel_palm"].reshape(N, -1),
        g["kp_rel_goal"].reshape(N, -1),
        st["object_scales"],
        st["closest_keypoint_max_dist"].unsqueeze(-1),
        st["closest_fingertip_dist"],
        st["lifted_object"].unsqueeze(-1).to(torch.float32),
        torch.log(st["progress_buf"] / 10 + 1).unsqueeze(-1),
        torch.log(st["successes"] + 1).unsqueeze(-1),
    ]
    body = torch.cat(parts, dim=-1)
    ofs = body.shape[1]
    st["obs_buf"][:, :ofs] = body
    return ofs


def observe(st: Dict[str, Tensor], c: KukaConsts) -> Dict[str, Tensor]:
    """compute_observations (allegro_kuka_base.py:991-1089): geometry, lazy init of closest/furthest, obs row."""
    g = _geometry(st, c)
    st["closest_fingertip_dist"] = torch.where(st["closest_fingertip_dist"] < 0.0, g["tip_dist"], st["closest_fingertip_dist"])
    st["furthest_hand_dist"] = torch.where(st["furthest_hand_dist"] < 0.0, g["tip_dist"][:, 0], st["furthest_hand_dist"])
    st["closest_keypoint_max_dist"] = torch.where(st["closest_keypoint_max_dist"] < 0.0, g["kp_max"], st["closest_keypoint_max_dist"])
    g["reward_ofs"] = _fill_obs(st, c, g)
    return g


def post_step(st: Dict[str, Tensor], c: KukaConsts) -> Dict[str, Tensor]:
    """One ``post_physics_step`` plus the VecTask.step tail; mutates ``st``; returns extras-like scalars."""
    st["progress_buf"] = st["progress_buf"] + 1
    st["randomize_buf"] = st["randomize_buf"] + 1
    g = observe(st, c)
    obj_z = g["obj"][:, 2]
    cur = g["tip_dist"]

    # lifting (:788-812)
    z_lift = 0.05 + obj_z - st["object_init_state"][:, 2]
    lifting_rew = torch.clip(z_lift, 0, 0.5)
    lifted_prev = st["lifted_object"]
    lifted = (z_lift > c.lifting_bonus_threshold) | lifted_prev
    lift_bonus_rew = c.lifting_bonus * (lifted & ~lifted_prev)
    lifting_rew = lifting_rew * ~lifted
    st["lifted_object"] = lifted

    # fingertip / hand distance deltas (:759-786)
    d_closest = st["closest_fingertip_dist"] - cur
    st["closest_fingertip_dist"] = torch.minimum(st["closest_fingertip_dist"], cur)
    d_furthest = st["furthest_hand_dist"] - cur[:, 0]
    st["furthest_hand_dist"] = torch.maximum(st["furthest_hand_dist"], cur[:, 0])
    ft = torch.sum(torch.clip(d_closest, 0, 10) * st["finger_rew_coeffs"], dim=-1)
    ft = ft * ~lifted
    hp = torch.clip(d_furthest, -10, 0)
    hp = hp * ~lifted
    hp = hp * c.n_tips

    # keypoints (:814-828)
    kp_delta = st["closest_keypoint_max_dist"] - g["kp_max"]
    st["closest_keypoint_max_dist"] = torch.minimum(st["closest_keypoint_max_dist"], g["kp_max"])
    kp_rew = torch.clip(kp_delta, 0, 100) * lifted

    # success logic (:857-869)
    near_goal = g["kp_max"] <= c.success_tolerance * c.keypoint_scale
    st["near_goal_steps"] = st["near_goal_steps"] + near_goal
    is_success = st["near_goal_steps"] >= c.success_steps
    st["successes"] = st["successes"] + is_success
    st["reset_goal_buf"][:] = is_success

    acc = st["rewards_episode"]
    acc[0] += ft
    acc[1] += hp
    acc[2] += lifting_rew
    acc[3] += kp_rew
    ft = ft * c.distance_delta_rew_scale
    hp = hp * (c.distance_delta_rew_scale * 0)
    lifting_rew = lifting_rew * c.lifting_rew_scale
    kp_rew = kp_rew * c.keypoint_rew_scale
    vel = st["dof_state"].view(cur.shape[0], -1, 2)[:, :c.n_dofs, 1]
    kuka_pen = -1 * (torch.sum(torch.abs(vel[:, 0:c.n_arm_dofs]), dim=-1) * c.kuka_actions_penalty_scale)
    allegro_pen = -1 * (torch.sum(torch.abs(vel[:, c.n_arm_dofs:c.n_dofs]), dim=-1) * c.allegro_actions_penalty_scale)
    bonus_rew = near_goal * (c.reach_goal_bonus / c.success_steps)
    reward = ft + hp + lifting_rew + lift_bonus_rew + kp_rew + kuka_pen + allegro_pen + bonus_rew
    st["rew_buf"][:] = reward

    # resets (:841-849, reorientation :97-102)
    ones = torch.ones_like(st["reset_buf"])
    resets = torch.where(obj_z < 0.1, ones, st["reset_buf"])
    if c.max_consecutive_successes > 0:
        st["progress_buf"] = torch.where(is_success, torch.zeros_like(st["progress_buf"]), st["progress_buf"])
        resets = torch.where(st["successes"] >= c.max_consecutive_successes, ones, resets)
    resets = torch.where(st["progress_buf"] >= c.max_episode_length - 1, ones, resets)
    if c.extra_far_reset and c.subtask == "reorientation":      # allegro_kuka_reorientation.py:97-102 only
        resets = torch.where(cur.max(dim=-1).values > 1.5, ones, resets)
    st["reset_buf"][:] = resets

    scale_, offset_ = true_objective_coeffs(c.success_tolerance, c.initial_tolerance, c.target_tolerance)
    st["true_objective"] = (st["successes"] * scale_) + offset_ if scale_ != 1.0 else st["successes"] + offset_

    step_terms = [ft, hp, lifting_rew, lift_bonus_rew, kp_rew, kuka_pen, allegro_pen, bonus_rew]
    for name, val in zip(STEP_TERMS, step_terms):
        acc[EPISODE_TERMS.index(name)] += val
        st["episode_cumulative"][STEP_TERMS.index(name)] = val

    # reward observation and clamp (:1437-1441)
    st["obs_buf"][:, g["reward_ofs"]] = reward * 0.01
    if c.clamp_abs_observations > 0:
        st["obs_buf"].clamp_(-c.clamp_abs_observations, c.clamp_abs_observations)
    # vec_task.py:424
    st["timeout_buf"] = (st["progress_buf"] >= c.max_episode_length - 1) & (st["reset_buf"] != 0)
    return {
        "successes_mean": st["prev_episode_successes"].mean(),
        "true_objective_mean": st["true_objective"].mean(),
        "true_objective_min": st["true_objective"].min(),
        "true_objective_max": st["true_objective"].max(),
        "is_success": is_success,
    }
