"""ORACLE (test infrastructure, not product code) -- CPU restatement of the two-arm AllegroKuka hot path (regrasping, reorientation).

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this.

Restates, op for op in the reference's association order,

* ``pre_step``  <- ``allegro_kuka_two_arms.py:1298-1351`` (pre_physics_step), ``:1215-1296`` (reset_idx), ``:1152-1157``
                   (reset_target_pose), ``:1159-1193`` (reset_object_pose, +-0.35 m side flip), ``:1139-1150`` (get_random_quat),
                   ``allegro_kuka_two_arms_regrasping.py:75-110`` (_reset_target: +-0.75 m side flip, the object is reset too),
                   ``allegro_kuka_two_arms_reorientation.py:97-137`` (_reset_target: side flip + random rotation, goal row velocities
                   zeroed, the object stays; four keypoints ``:25-31``, 196-wide row)
* ``post_step`` <- ``allegro_kuka_two_arms.py:1353-1370`` (post_physics_step), ``:939-1133`` (observations, 169-wide row),
                   ``:743-878`` (rewards: no finger coefficients, no lifted gate on the fingertip term, no hand / action penalties),
                   ``allegro_kuka_utils.py:128-158`` (true objective), ``vec_task.py:424`` (timeout_buf)

Pinned against the live reference by ``tests/test_kuka2_oracle_vs_reference.py`` and the reference-generated
``tests/golden/kuka2_*.npz``.  Random numbers: per-env table laid out as ``HA_KUKA2_RND_*``.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Tuple

import numpy as np
import torch

from .kuka_oracle import quat_rotate, random_quat, scale, true_objective_coeffs, unscale
from .kuka_state import KUKA_LOWER, KUKA_UPPER

Tensor = torch.Tensor

# per-env random table (include/handarm_b200.h, HA_KUKA2_RND_*)
RND_GOAL, RND_RESET_GOAL, RND_RESET_OBJ, RND_FORCE_PROB, RND_FORCE_TEST, RND_FORCE_NORMAL, RND_DOF_POS = 0, 11, 22, 29, 30, 31, 34
# inside a goal block: target position 3, target side 1, then (regrasping) object side 1, object position 3, object rotation 3
# or (reorientation) target rotation uvw 3
G_POS, G_SIDE, G_OBJ, G_ROT = 0, 3, 4, 4
# inside an object block: side 1, position 3, rotation 3
O_SIDE, O_POS, O_ROT = 0, 1, 4

EPISODE_TERMS = ["raw_fingertip_delta_rew", "raw_lifting_rew", "raw_keypoint_rew", "fingertip_delta_rew", "lifting_rew", "lift_bonus_rew",
                 "keypoint_rew", "bonus_rew"]                                       # allegro_kuka_two_arms.py:345-354
STEP_TERMS = ["fingertip_delta_rew", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "bonus_rew"]   # :861-867


def rnd_stride(n_dofs: int = 46) -> int:
    return RND_DOF_POS + 2 * n_dofs


@dataclasses.dataclass
class Kuka2Consts:
    n_arms: int = 2
    n_arm_dofs: int = 7
    n_dofs: int = 46
    n_tips: int = 8
    palm_bodies: Tuple[int, ...] = (7, 31)
    tip_bodies: Tuple[int, ...] = (11, 15, 19, 23, 35, 39, 43, 47)
    object_body: int = 48
    n_bodies: int = 51
    n_actors: int = 5                         # arm0, arm1, object, table, goal ball
    palm_offset: Tuple[float, ...] = (-0.0, -0.02, 0.16)
    tip_offsets: Tuple[Tuple[float, ...], ...] = ((0.05, 0.005, 0), (0.05, 0.005, 0), (0.05, 0.005, 0), (0.06, 0.005, 0))
    dt: float = 0.01667
    clamp_abs_observations: float = 10.0
    max_episode_length: int = 300             # env/regrasping.yaml
    max_consecutive_successes: int = 50
    success_steps: int = 30
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
    target_volume_origin: Tuple[float, ...] = (0, 0.0, 0.8)
    target_volume_extent: Tuple[Tuple[float, float], ...] = ((-0.2, 0.2), (-0.5, 0.5), (-0.12, 0.25))
    goal_x_ofs: float = 0.75
    obj_x_ofs: float = 1.1 / 2 - 0.2
    subtask: str = "regrasping"               # or "reorientation" (allegro_kuka_two_arms_reorientation.py)

    @property
    def n_keypoints(self) -> int:
        return 4 if self.subtask == "reorientation" else 1

    @property
    def obs_dim(self) -> int:
        return 169 + 9 * (self.n_keypoints - 1)   # K = 1: 169, K = 4: 196 (allegro_kuka_two_arms.py:176-211)


# ---------------------------------------------------------------------------------------------
# pre-physics step
# ---------------------------------------------------------------------------------------------
def _reset_object_pose(st, c: Kuka2Consts, ids: Tensor, blk: Tensor) -> None:
    """allegro_kuka_two_arms.py:1159-1193; ``blk`` = raw draws [n, 7]: side, position 3, rotation 3."""
    rows = st["object_indices"][ids]
    left_right = 2.0 * blk[:, O_SIDE:O_SIDE + 1] + -1.0
    x_pos = torch.where(left_right > 0, c.obj_x_ofs * torch.ones_like(left_right), -c.obj_x_ofs * torch.ones_like(left_right))
    r = 2.0 * blk[:, O_POS:O_POS + 3] + -1.0
    init = st["object_init_state"][ids]
    st["root_state"][rows] = init.clone()
    st["root_state"][rows, 0:1] = x_pos + c.reset_position_noise[0] * r[:, 0:1]
    st["root_state"][rows, 1:2] = init[:, 1:2] + c.reset_position_noise[1] * r[:, 1:2]
    st["root_state"][rows, 2:3] = init[:, 2:3] + c.reset_position_noise[2] * r[:, 2:3]
    st["root_state"][rows, 3:7] = random_quat(blk[:, O_ROT:O_ROT + 3])
    st["root_state"][rows, 7:13] = 0.0
    st["closest_fingertip_dist"][ids] = -1.0


def _reset_target_pose(st, c: Kuka2Consts, ids: Tensor, blk: Tensor) -> None:
    """reset_target_pose (:1152-1157) with the subtask's ``_reset_target`` (allegro_kuka_two_arms_regrasping.py:75-110,
    allegro_kuka_two_arms_reorientation.py:97-137)."""
    org = torch.tensor(c.target_volume_origin, dtype=torch.float32)
    ext = torch.tensor(c.target_volume_extent, dtype=torch.float32)
    lo = org + ext[:, 0]
    size = (org + ext[:, 1]) - lo
    coords = lo + blk[:, G_POS:G_POS + 3] * size
    left_right = 2.0 * blk[:, G_SIDE:G_SIDE + 1] + -1.0
    x_pos = torch.where(left_right > 0, c.goal_x_ofs * torch.ones_like(left_right), -c.goal_x_ofs * torch.ones_like(left_right))
    coords[:, 0] += x_pos.squeeze(dim=1)
    st["goal_states"][ids, 0:3] = coords
    rows = st["goal_object_indices"][ids]
    st["root_state"][rows, 0:3] = st["goal_states"][ids, 0:3]
    if c.subtask == "reorientation":          # allegro_kuka_two_arms_reorientation.py:125-134
        st["goal_states"][ids, 3:7] = random_quat(blk[:, G_ROT:G_ROT + 3])
        st["root_state"][rows, 3:7] = st["goal_states"][ids, 3:7]
        st["root_state"][rows, 7:13] = 0.0
    else:                                     # allegro_kuka_two_arms_regrasping.py:100-106
        _reset_object_pose(st, c, ids, blk[:, G_OBJ:G_OBJ + 7])
        st["lifted_object"][ids] = False
    st["reset_goal_buf"][ids] = 0
    st["near_goal_steps"][ids] = 0
    st["closest_keypoint_max_dist"][ids] = -1.0


def pre_step(st: Dict[str, Tensor], c: Kuka2Consts, actions: Tensor, rnd: Tensor) -> Dict[str, Tensor]:
    """allegro_kuka_two_arms.py:1298-1351.  Mutates ``st``; returns the actor-row lists handed to the simulator."""
    D = c.n_dofs
    lo, hi = st["dof_lower"], st["dof_upper"]
    st["actions"] = actions.clone()
    ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    root_lists: List[Tensor] = []
    moves_object = c.subtask != "reorientation"      # regrasping's _reset_target defers the object row as well
    _reset_target_pose(st, c, goal_ids, rnd[goal_ids, RND_GOAL:RND_GOAL + 11])
    root_lists += ([st["object_indices"][goal_ids]] if moves_object else []) + [st["goal_object_indices"][goal_ids]]
    hand_rows = torch.zeros(0, dtype=torch.int32)
    if len(ids) > 0:
        _reset_target_pose(st, c, ids, rnd[ids, RND_RESET_GOAL:RND_RESET_GOAL + 11])
        root_lists += ([st["object_indices"][ids]] if moves_object else []) + [st["goal_object_indices"][ids]]
        st["rb_forces"][ids] = 0.0
        _reset_object_pose(st, c, ids, rnd[ids, RND_RESET_OBJ:RND_RESET_OBJ + 7])
        hand_rows = st["arm_indices"][ids].to(torch.int32).flatten()
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        st["random_force_prob"][ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, RND_FORCE_PROB] + torch.log(pr[1]))
        dflt, coeff = st["dof_default"], st["pos_noise_coeff"]
        delta_max, delta_min = hi - dflt, lo - dflt
        rand_delta = delta_min + (delta_max - delta_min) * rnd[ids, RND_DOF_POS:RND_DOF_POS + D]
        pos = dflt + coeff * rand_delta
        st["dof_state"][ids, :D, 0] = pos
        st["prev_targets"][ids, :D] = pos
        st["cur_targets"][ids, :D] = pos
        st["dof_state"][ids, :D, 1] = c.reset_dof_vel_noise * (2.0 * rnd[ids, RND_DOF_POS + D:RND_DOF_POS + 2 * D] + -1.0)
        root_lists += [st["object_indices"][ids], st["goal_object_indices"][ids]]
        st["progress_buf"][ids] = 0
        st["reset_buf"][ids] = 0
        st["prev_episode_successes"][ids] = st["successes"][ids]
        st["successes"][ids] = 0.0
        st["prev_episode_true_objective"][ids] = st["true_objective"][ids]
        st["true_objective"][ids] = 0.0
        st["lifted_object"][ids] = False
        st["closest_keypoint_max_dist"][ids] = -1.0
        st["closest_fingertip_dist"][ids] = -1.0
        st["near_goal_steps"][ids] = 0
        st["rewards_episode"][:, ids] = 0.0
    root_rows = torch.unique(torch.cat(root_lists).to(torch.int32))
    cur, prev = st["cur_targets"], st["prev_targets"]
    t = scale(actions[..., :D], lo[:D], hi[:D])
    t = c.act_moving_average * t + (1.0 - c.act_moving_average) * prev[..., :D]
    cur[..., :D] = torch.max(torch.min(t, hi[:D]), lo[:D])
    prev[...] = cur[...]
    if c.force_scale > 0.0:
        st["rb_forces"] *= torch.pow(torch.tensor(c.force_decay, dtype=torch.float32), c.dt / c.force_decay_interval)
        hit = (rnd[:, RND_FORCE_TEST] < st["random_force_prob"]).nonzero(as_tuple=False).squeeze(-1)
        st["rb_forces"][hit, c.object_body, :] = rnd[hit, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] * c.object_mass * c.force_scale
    return {"root_rows": root_rows, "hand_rows": hand_rows}


# ---------------------------------------------------------------------------------------------
# post-physics step
# ---------------------------------------------------------------------------------------------
def _geometry(st, c: Kuka2Consts) -> Dict[str, Tensor]:
    """allegro_kuka_two_arms.py:944-1025."""
    N, A, F = st["reset_buf"].shape[0], c.n_arms, c.n_tips // c.n_arms
    obj = st["root_state"][st["object_indices"]]
    rb = st["rigid_body_state"].view(N, -1, 13)
    palm = rb[:, list(c.palm_bodies)]                                  # [N, A, 13]
    tips = rb[:, list(c.tip_bodies)][:, :, 0:13]
    g: Dict[str, Tensor] = {"obj": obj, "palm": palm}
    palm_off = torch.tensor(c.palm_offset, dtype=torch.float32).repeat(N, 1)
    pc = torch.zeros(N, A, 3)
    for a in range(A):
        pc[:, a] = palm[:, a, 0:3] + quat_rotate(palm[:, a, 3:7], palm_off)
    g["palm_center"] = pc
    tip_off = torch.tensor(np.asarray(c.tip_offsets, dtype=np.float32)).repeat(N, 1, 1)
    tip_pos = torch.zeros(N, A * F, 3)
    for a in range(A):
        for i in range(F):
            f = a * F + i
            tip_pos[:, f] = tips[:, f, 0:3] + quat_rotate(tips[:, f, 3:7], tip_off[:, i])
    rel_obj = tip_pos - obj[:, None, 0:3]
    g["tip_dist"] = torch.norm(rel_obj.view(N, A, F, -1), dim=-1)      # [N, A, F]
    g["tip_rel_palm"] = tip_pos - pc.unsqueeze(2).repeat(1, 1, F, 1).view(N, A * F, 3)
    K = st["object_keypoint_offsets"].shape[1]
    okp, gkp = torch.zeros(N, K, 3), torch.zeros(N, K, 3)
    goal = st["goal_states"]
    for k in range(K):
        okp[:, k] = obj[:, 0:3] + quat_rotate(obj[:, 3:7], st["object_keypoint_offsets"][:, k])
        gkp[:, k] = goal[:, 0:3] + quat_rotate(goal[:, 3:7], st["object_keypoint_offsets"][:, k])
    g["kp_rel_goal"] = okp - gkp
    g["kp_rel_palm"] = (okp.unsqueeze(1).repeat(1, A, 1, 1) - pc.unsqueeze(2).repeat(1, 1, K, 1)).view(N, A * K, 3)
    g["kp_max"] = torch.norm(g["kp_rel_goal"], dim=-1).max(dim=-1).values
    return g


def observe(st, c: Kuka2Consts) -> Dict[str, Tensor]:
    """compute_observations + compute_full_state (allegro_kuka_two_arms.py:939-1133)."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    g = _geometry(st, c)
    st["closest_fingertip_dist"] = torch.where(st["closest_fingertip_dist"] < 0.0, g["tip_dist"], st["closest_fingertip_dist"])
    st["closest_keypoint_max_dist"] = torch.where(st["closest_keypoint_max_dist"] < 0.0, g["kp_max"], st["closest_keypoint_max_dist"])
    dof = st["dof_state"].view(N, -1, 2)[:, :D]
    body = torch.cat([
        unscale(dof[..., 0], st["dof_lower"][:D], st["dof_upper"][:D]), dof[..., 1], g["palm_center"].view(N, -1),
        g["palm"][..., 3:13].reshape(N, -1), g["obj"][:, 3:13], g["tip_rel_palm"].reshape(N, -1), g["kp_rel_palm"].reshape(N, -1),
        g["kp_rel_goal"].reshape(N, -1), st["object_scales"], st["closest_keypoint_max_dist"].unsqueeze(-1),
        st["lifted_object"].unsqueeze(-1).to(torch.float32), torch.log(st["progress_buf"] / 10 + 1).unsqueeze(-1),
        torch.log(st["successes"] + 1).unsqueeze(-1)], dim=-1)
    g["reward_ofs"] = body.shape[1]
    st["obs_buf"][:, : body.shape[1]] = body
    return g


def post_step(st, c: Kuka2Consts) -> Dict[str, Tensor]:
    """allegro_kuka_two_arms.py:1353-1370 plus the VecTask.step tail."""
    st["progress_buf"] = st["progress_buf"] + 1
    st["randomize_buf"] = st["randomize_buf"] + 1
    g = observe(st, c)
    obj_z = g["obj"][:, 2]
    cur = g["tip_dist"]
    z_lift = 0.05 + obj_z - st["object_init_state"][:, 2]
    lifting_rew = torch.clip(z_lift, 0, 0.5)
    lifted_prev = st["lifted_object"]
    lifted = (z_lift > c.lifting_bonus_threshold) | lifted_prev
    lift_bonus_rew = c.lifting_bonus * (lifted & ~lifted_prev)
    lifting_rew = lifting_rew * ~lifted
    st["lifted_object"] = lifted
    d_closest = st["closest_fingertip_dist"] - cur
    st["closest_fingertip_dist"] = torch.minimum(st["closest_fingertip_dist"], cur)
    ft = torch.sum(torch.sum(torch.clip(d_closest, 0, 10), dim=-1), dim=-1)
    kp_delta = st["closest_keypoint_max_dist"] - g["kp_max"]
    st["closest_keypoint_max_dist"] = torch.minimum(st["closest_keypoint_max_dist"], g["kp_max"])
    kp_rew = torch.clip(kp_delta, 0, 100) * lifted
    near_goal = g["kp_max"] <= c.success_tolerance * c.keypoint_scale
    st["near_goal_steps"] = st["near_goal_steps"] + near_goal
    is_success = st["near_goal_steps"] >= c.success_steps
    st["successes"] = st["successes"] + is_success
    st["reset_goal_buf"][:] = is_success
    acc = st["rewards_episode"]
    acc[0] += ft
    acc[1] += lifting_rew
    acc[2] += kp_rew
    ft = ft * c.distance_delta_rew_scale
    lifting_rew = lifting_rew * c.lifting_rew_scale
    kp_rew = kp_rew * c.keypoint_rew_scale
    bonus_rew = near_goal * (c.reach_goal_bonus / c.success_steps)
    reward = ft + lifting_rew + lift_bonus_rew + kp_rew + bonus_rew
    st["rew_buf"][:] = reward
    ones = torch.ones_like(st["reset_buf"])
    resets = torch.where(obj_z < 0.1, ones, st["reset_buf"])
    if c.max_consecutive_successes > 0:
        st["progress_buf"] = torch.where(is_success > 0, torch.zeros_like(st["progress_buf"]), st["progress_buf"])
        resets = torch.where(st["successes"] >= c.max_consecutive_successes, ones, resets)
    resets = torch.where(st["progress_buf"] >= c.max_episode_length - 1, ones, resets)
    st["reset_buf"][:] = resets
    scale_, offset_ = true_objective_coeffs(c.success_tolerance, c.initial_tolerance, c.target_tolerance)
    st["true_objective"] = (st["successes"] * scale_) + offset_ if scale_ != 1.0 else st["successes"] + offset_
    for name, val in zip(STEP_TERMS, [ft, lifting_rew, lift_bonus_rew, kp_rew, bonus_rew]):
        acc[EPISODE_TERMS.index(name)] += val
        st["episode_cumulative"][STEP_TERMS.index(name)] = val
    st["obs_buf"][:, g["reward_ofs"]] = reward * 0.01
    if c.clamp_abs_observations > 0:
        st["obs_buf"].clamp_(-c.clamp_abs_observations, c.clamp_abs_observations)
    st["timeout_buf"] = (st["progress_buf"] >= c.max_episode_length - 1) & (st["reset_buf"] != 0)
    return {"successes_mean": st["prev_episode_successes"].mean(), "true_objective_mean": st["true_objective"].mean(),
            "true_objective_min": st["true_objective"].min(), "true_objective_max": st["true_objective"].max(), "is_success": is_success}


# ---------------------------------------------------------------------------------------------
# seeded states
# ---------------------------------------------------------------------------------------------
def make_state(n: int, seed: int, c: Kuka2Consts = Kuka2Consts(), fresh: bool = False) -> Dict[str, Tensor]:
    gen = torch.Generator().manual_seed(seed)
    D, T, B, A = c.n_dofs, c.n_tips, c.n_bodies, c.n_actors
    U = lambda *s: torch.rand(*s, generator=gen)
    G = lambda *s: torch.randn(*s, generator=gen)
    unit = lambda *s: (lambda q: q / q.norm(dim=-1, keepdim=True).clamp_min(1e-6))(G(*s, 4))
    st: Dict[str, Tensor] = {}
    ar = torch.arange(n, dtype=torch.long) * A
    st["arm_indices"] = torch.stack([ar + 0, ar + 1], dim=1)
    st["object_indices"], st["goal_object_indices"] = ar + 2, ar + 4
    st["dof_lower"] = torch.tensor(KUKA_LOWER * 2, dtype=torch.float32)
    st["dof_upper"] = torch.tensor(KUKA_UPPER * 2, dtype=torch.float32)
    dflt = torch.zeros(2, 23)
    dflt[:, :7] = torch.tensor([-1.571, 1.571, -0.000, 1.6, -0.000, 1.485, 2.358])
    st["dof_default"] = dflt.flatten()
    coeff = torch.zeros(2, 23)
    coeff[:, 0:7] = c.reset_dof_pos_noise_arm
    coeff[:, 7:23] = c.reset_dof_pos_noise_fingers
    st["pos_noise_coeff"] = coeff.flatten()
    init = torch.zeros(n, 13)
    init[:, 0:3] = torch.tensor([0.0, 0.0, 0.63])
    init[:, 6] = 1.0
    st["object_init_state"] = init
    goal = init.clone()
    goal[:, 0:3] = torch.tensor([0.0, 0.0, 0.8]) + 0.2 * G(n, 3)
    st["goal_states"] = goal
    scales = torch.tensor([[1.0, 1.0, 1.0], [4.0, 0.75, 1.0], [1.0, 2.5, 1.0], [1.25, 0.66, 3.0], [0.5, 2.0, 1.5]])
    st["object_scales"] = scales[torch.arange(n) % len(scales)].contiguous()
    if c.subtask == "reorientation":          # four corners (allegro_kuka_two_arms_reorientation.py:25-31), scaled (:692-700)
        corners = torch.tensor([[1, 1, 1], [1, 1, -1], [-1, -1, 1], [-1, -1, -1]], dtype=torch.float32)
        st["object_keypoint_offsets"] = (corners[None] * (st["object_scales"][:, None, :] * 0.05 * c.keypoint_scale / 2)).contiguous()
        goal[:, 3:7] = unit(n)
    else:
        st["object_keypoint_offsets"] = torch.zeros(n, 1, 3)            # regrasping: one keypoint at the object centre
    root = torch.zeros(n, A, 13)
    root[..., 0:3] = torch.tensor([0.0, 0.0, 0.63]) + 0.15 * G(n, A, 3)
    root[..., 3:7] = unit(n, A)
    root[..., 7:10] = 0.5 * G(n, A, 3)
    root[..., 10:13] = 2.0 * G(n, A, 3)
    root[U(n) < 0.04, 2, 2] = 0.05
    hit = U(n) < 0.15
    root[hit, 2, 0:3] = goal[hit, 0:3] + 1e-2 * G(int(hit.sum()), 3)
    rb = torch.zeros(n, B, 13)
    rb[..., 0:3] = root[:, 2:3, 0:3] + 0.15 * G(n, B, 3)
    rb[..., 3:7] = unit(n, B)
    rb[..., 7:10] = 0.5 * G(n, B, 3)
    rb[..., 10:13] = 2.0 * G(n, B, 3)
    rb[:, c.object_body] = root[:, 2]
    st["root_state"] = root.view(-1, 13).contiguous()
    st["rigid_body_state"] = rb.view(-1, 13).contiguous()
    dof = torch.zeros(n, D, 2)
    dof[..., 0] = st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)
    dof[..., 1] = G(n, D)
    st["dof_state"] = dof.contiguous()
    st["obs_buf"] = torch.zeros(n, c.obs_dim)
    st["rew_buf"] = torch.zeros(n)
    st["timeout_buf"] = torch.zeros(n, dtype=torch.bool)
    st["actions"] = torch.zeros(n, D)
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * U(n) + torch.log(pr[1]))
    st["episode_cumulative"] = torch.zeros(len(STEP_TERMS), n)
    if fresh:
        st.update(reset_buf=torch.ones(n, dtype=torch.long), reset_goal_buf=torch.ones(n, dtype=torch.long),
                  progress_buf=torch.zeros(n, dtype=torch.long), randomize_buf=torch.zeros(n, dtype=torch.long), successes=torch.zeros(n),
                  prev_episode_successes=torch.zeros(n), true_objective=torch.zeros(n), prev_episode_true_objective=torch.zeros(n),
                  near_goal_steps=torch.zeros(n, dtype=torch.int32), lifted_object=torch.zeros(n, dtype=torch.bool),
                  closest_keypoint_max_dist=-torch.ones(n), closest_fingertip_dist=-torch.ones(n, 2, 4), prev_targets=torch.zeros(n, D),
                  cur_targets=torch.zeros(n, D), rb_forces=torch.zeros(n, B, 3), rewards_episode=torch.zeros(len(EPISODE_TERMS), n))
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
    near = torch.randint(0, max(c.success_steps, 1), (n,), generator=gen).to(torch.int32)
    near[U(n) < 0.2] = max(c.success_steps - 1, 0)
    st["near_goal_steps"] = near
    st["lifted_object"] = U(n) < 0.4
    ckp = 0.05 + 0.3 * U(n)
    ckp[U(n) < 0.2] = -1.0
    st["closest_keypoint_max_dist"] = ckp
    ctip = 0.05 + 0.3 * U(n, 2, 4)
    ctip[U(n) < 0.2] = -1.0
    st["closest_fingertip_dist"] = ctip
    st["prev_targets"] = (st["dof_lower"] + (st["dof_upper"] - st["dof_lower"]) * U(n, D)).contiguous()
    st["cur_targets"] = st["prev_targets"].clone()
    frc = torch.zeros(n, B, 3)
    frc[:, c.object_body] = 0.1 * G(n, 3) * (U(n, 1) < 0.3)
    st["rb_forces"] = frc
    st["rewards_episode"] = G(len(EPISODE_TERMS), n).contiguous()
    return st


def make_actions(n: int, seed: int, c: Kuka2Consts = Kuka2Consts()) -> Tensor:
    gen = torch.Generator().manual_seed(seed + 7919)
    return (2.0 * torch.rand(n, c.n_dofs, generator=gen) - 1.0).contiguous()


def make_rnd(n: int, seed: int, c: Kuka2Consts = Kuka2Consts()) -> Tensor:
    gen = torch.Generator().manual_seed(seed + 104729)
    rnd = torch.rand(n, rnd_stride(c.n_dofs), generator=gen)
    rnd[:, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] = torch.randn(n, 3, generator=gen)
    return rnd.contiguous()
