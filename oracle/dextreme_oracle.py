"""ORACLE (test infrastructure, not product code) -- CPU restatement of the AllegroHandDextremeADR per-step tensor path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this.

Torch-on-CPU restatement, op for op in the reference's association order, of

* ``pre_step``  <- ``allegro_hand_dextreme.py:912-949`` (pre_physics_step with the ADR ``apply_reset_buf`` one-step delay),
                   ``:709-733`` (reset_target_pose), ``:775-858`` (reset_idx), ``:695-707`` (get_random_quat),
                   ``:955-1015`` (apply_actions: action queue, control, moving average), ``:1430-1441`` (ADR
                   apply_action_noise_latency), ``:1017-1030`` (apply_random_forces)
* ``post_step`` <- ``:1033-1048`` (post_physics_step), ``:606-692`` + ``:1443-1489`` (compute_observations, base + ADR),
                   ``:533-588`` (compute_poses_wrt_wrist), ``:739-771`` (get_random_cube_observation),
                   ``:1383-1387`` (sample_gaussian_adr), ``:438-460`` + ``:1598-1663`` (compute_reward / compute_hand_reward),
                   ``adr_vec_task.py:289-325`` (step tail: timeout_buf, clamp of every dict entry),
                   ``torch_jit_utils.py:462-544`` (quaternion_to_matrix / matrix_to_quaternion, wxyz)

Outside this restatement (host control logic): ``ADRVecTask.apply_randomizations`` (physics-parameter DR through the
Isaac Gym API, ADR range update, per-env ADR resampling) and the RNA network -- the per-env ADR tensors and affine parameters are
inputs here, and an optional ``rna_perturbation`` [N,16] (already unscaled) is blended with ``rna_alpha`` as ``:887-890`` does.
The reference's relative-control branch (``:978-981``) reads an attribute that is never defined and cannot run; absolute and
capped control are restated.

Parity pinning: ``tests/test_dextreme_oracle_vs_reference.py`` runs the reference's real ``AllegroHandDextremeADR`` methods (through
``oracle/reference_loader.py``, with the one documented ``torch.where(goal_reached != 0, ...)`` patch at ``:1630`` that torch >= 2
needs) and this restatement on the same inputs; ``tests/golden/dextreme_*.npz`` hold reference-generated vectors.  Random numbers
come from two per-env tables laid out as ``HA_DEX_RND_*`` (pre) and ``HA_DEX_POST_RND_*`` (post).
"""
from __future__ import annotations

import dataclasses
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from .hand_oracle import quat_conjugate, quat_mul, randomize_rotation
from .kuka_oracle import random_quat, scale, unscale

Tensor = torch.Tensor

# ---- pre-step random table (raw U(0,1) except *_NORMAL / *_WHITE slots, which hold N(0,1) draws) ----
RND_GOAL = 0            # 4 rand_floats + 3 quaternion uniforms of a goal-only reset
RND_RESET_GOAL = 7      # the same 7 for reset_target_pose called from reset_idx
RND_RESET = 14          # 2*D+5 rand_floats + 3 quaternion uniforms
RND_FORCE_PROB = 54
RND_PROGRESS = 55       # stands in for torch.randint(0, max_episode_length) on an all-env reset
RND_ACTION_DELAY = 56
RND_ACTION_WHITE = 57   # 16 normals
RND_FORCE_TEST = 73
RND_FORCE_NORMAL = 74   # 3 normals
RND_STRIDE = 77
# ---- post-step random table ----
PRND_POSE_WHITE = 0     # 7 normals
PRND_CUBE = 7           # 5 rand_floats
PRND_CUBE_QUAT = 12     # 3 uniforms
PRND_CUBE_MASK = 15
PRND_DELAY = 16
PRND_DOF_WHITE = 17     # 16 normals
PRND_STRIDE = 33

OBS_KEYS: List[Tuple[str, int]] = [
    ("dof_pos", 16), ("dof_pos_randomized", 16), ("dof_vel", 16), ("dof_force", 16), ("object_vels", 6), ("last_actions", 16),
    ("cube_random_params", 3), ("hand_random_params", 1), ("gravity_vec", 3), ("ft_states", 52), ("ft_force_torques", 24),
    ("rb_forces", 3), ("rot_dist", 2), ("stochastic_delay_params", 4), ("affine_params", 78), ("object_pose", 7), ("goal_pose", 7),
    ("goal_relative_rot", 4), ("object_pose_cam_randomized", 7), ("goal_relative_rot_cam_randomized", 4)]      # :168-199
OBS_DIM = sum(k for _, k in OBS_KEYS)                                                                         # 285
STEP_TERMS = ["dist_rew", "rot_rew", "action_penalty", "action_delta_penalty", "velocity_penalty", "reach_goal_rew", "fall_rew",
              "timeout_rew"]
ADR_KEYS = ["action_delay_prob", "affine_action_white", "affine_cube_pose_white", "cube_obs_delay_prob", "affine_dof_pos_white",
            "rna_alpha"]


def obs_offsets() -> Dict[str, Tuple[int, int]]:
    out, o = {}, 0
    for k, w in OBS_KEYS:
        out[k] = (o, w)
        o += w
    return out


@dataclasses.dataclass
class DexConsts:
    n_dofs: int = 16
    n_tips: int = 4
    n_bodies: int = 19
    palm_body: int = 0
    tip_bodies: Tuple[int, ...] = (4, 8, 12, 16)
    object_body: int = 17
    dt: float = 0.01667
    control_freq_inv: int = 2
    max_episode_length: int = 240                    # round(resetTime 8 / (2 * 0.01667))
    clip_obs: float = 50.0
    latency_max: int = 15                            # actionLatencyMax -> queue of 16 slots
    use_capped_dof_control: bool = False
    max_dof_radians_per_second: float = 6.2832
    act_moving_average: float = 0.2                  # upper end of the annealing range at step 0
    apply_random_quat: bool = True
    random_cube_pose_prob: float = 0.3
    dist_reward_scale: float = -10.0
    rot_reward_scale: float = 1.0
    rot_eps: float = 0.1
    action_penalty_scale: float = -0.001
    action_delta_penalty_scale: float = -0.2
    success_tolerance: float = 0.1
    reach_goal_bonus: float = 250.0
    fall_dist: float = 0.24
    fall_penalty: float = 0.0
    max_consecutive_successes: int = 50
    av_factor: float = 0.1
    num_success_hold_steps: int = 0
    vel_obs_scale: float = 1.0                       # ADR subclass overrides both scales to 1 (:1494-1495)
    force_torque_obs_scale: float = 1.0
    reset_position_noise: float = 0.03
    reset_position_noise_z: float = 0.01
    reset_dof_pos_noise: float = 0.2
    reset_dof_vel_noise: float = 0.0
    force_scale: float = 2.0
    force_decay: float = 0.99
    force_decay_interval: float = 0.08
    force_prob_range: Tuple[float, float] = (0.001, 0.1)
    object_mass: float = 0.05
    goal_displacement: Tuple[float, float, float] = (-0.2, -0.06, 0.12)


def gaussian_stdev(adr_value: Tensor) -> Tensor:
    """sample_gaussian_adr's nonlinearity (:1383-1387): exp(v^2) - 1 where v > 0, else 0; [N] -> [N,1]."""
    v = adr_value.view(-1, 1)
    return torch.where(v > 0, torch.exp(torch.pow(v, 2.)) - 1., torch.zeros_like(v))


def dex_random_quat(uvw: Tensor) -> Tensor:
    """get_random_quat (:695-707) -- the same three-uniform construction as the Kuka tasks."""
    return random_quat(uvw)


# ---------------------------------------------------------------------------------------------
# wrist-relative poses (:533-588) on torch's own inverse / matmul, wxyz conversions as torch_jit_utils.py:462-544
# ---------------------------------------------------------------------------------------------
def quaternion_to_matrix(q: Tensor) -> Tensor:
    r, i, j, k = torch.unbind(q, -1)
    two_s = 2.0 / (q * q).sum(-1)
    m = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return m.reshape(q.shape[:-1] + (3, 3))


def matrix_to_quaternion(matrix: Tensor) -> Tensor:
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = torch.unbind(matrix.reshape(matrix.shape[:-2] + (9,)), dim=-1)
    x = torch.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22, 1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], dim=-1)
    q_abs = torch.zeros_like(x)
    q_abs[x > 0] = torch.sqrt(x[x > 0])
    cand = torch.stack([
        torch.stack([q_abs[..., 0] ** 2, m21 - m12, m02 - m20, m10 - m01], dim=-1),
        torch.stack([m21 - m12, q_abs[..., 1] ** 2, m10 + m01, m02 + m20], dim=-1),
        torch.stack([m02 - m20, m10 + m01, q_abs[..., 2] ** 2, m12 + m21], dim=-1),
        torch.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[..., 3] ** 2], dim=-1)], dim=-2)
    cand = cand / (2.0 * q_abs[..., None].max(torch.tensor(0.1)))
    pick = torch.nn.functional.one_hot(q_abs.argmax(dim=-1), num_classes=4) > 0.5
    return cand[pick, :].reshape(matrix.shape[:-2] + (4,))


def _pose_matrix(pose: Tensor) -> Tensor:
    T = torch.eye(4).repeat(pose.shape[0], 1, 1)
    T[:, 0:3, 0:3] = quaternion_to_matrix(pose[:, 3:7][:, [3, 0, 1, 2]])
    T[:, 0:3, 3] = pose[:, 0:3]
    return T


def pose_wrt_wrist(pose: Tensor, palm_pose: Tensor) -> Tensor:
    rel = torch.matmul(torch.inverse(_pose_matrix(palm_pose)), _pose_matrix(pose))
    return torch.cat((rel[:, 0:3, 3], matrix_to_quaternion(rel[:, 0:3, 0:3])[:, [1, 2, 3, 0]]), dim=-1)


# ---------------------------------------------------------------------------------------------
# pre-physics step
# ---------------------------------------------------------------------------------------------
def _new_rot(c: DexConsts, r0: Tensor, r1: Tensor, uvw: Tensor) -> Tensor:
    return dex_random_quat(uvw) if c.apply_random_quat else randomize_rotation(r0, r1)


def _reset_target_pose(st, c: DexConsts, ids: Tensor, blk: Tensor) -> None:
    """:709-733; ``blk`` [n,7] = 4 raw rand_floats + 3 quaternion uniforms."""
    rf = 2.0 * blk[:, 0:4] + -1.0
    rot = _new_rot(c, rf[:, 0], rf[:, 1], blk[:, 4:7])
    rows = st["goal_object_indices"][ids]
    st["goal_states"][ids, 0:3] = st["goal_init_state"][ids, 0:3]
    st["goal_states"][ids, 3:7] = rot
    st["root_state"][rows, 0:3] = st["goal_states"][ids, 0:3] + torch.tensor(c.goal_displacement, dtype=torch.float32)
    st["root_state"][rows, 3:7] = st["goal_states"][ids, 3:7]
    st["root_state"][rows, 7:13] = 0.0
    st["reset_goal_buf"][ids] = 0
    st["best_rotation_dist"][ids] = -1.0


def _reset_idx(st, c: DexConsts, ids: Tensor, rnd: Tensor) -> None:
    """:775-858."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    rf = 2.0 * rnd[ids, RND_RESET:RND_RESET + 2 * D + 5] + -1.0
    _reset_target_pose(st, c, ids, rnd[ids, RND_RESET_GOAL:RND_RESET_GOAL + 7])
    st["rb_forces"][ids, :, :] = 0.0
    rows = st["object_indices"][ids]
    init = st["object_init_state"][ids]
    st["root_state"][rows] = init.clone()
    st["root_state"][rows, 0:2] = init[:, 0:2] + c.reset_position_noise * rf[:, 0:2]
    st["root_state"][rows, 2] = init[:, 2] + c.reset_position_noise_z * rf[:, 2]
    st["root_state"][rows, 3:7] = _new_rot(c, rf[:, 3], rf[:, 4], rnd[ids, RND_RESET + 2 * D + 5:RND_RESET + 2 * D + 8])
    st["root_state"][rows, 7:13] = 0.0
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"][ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, RND_FORCE_PROB] + torch.log(pr[1]))
    lo, hi, dflt = st["dof_lower"], st["dof_upper"], st["dof_default"]
    delta_max, delta_min = hi - dflt, lo - dflt
    rand_delta = delta_min + (delta_max - delta_min) * ((rf[:, 5:5 + D] + 1) / 2)
    pos = torch.zeros(D) + c.reset_dof_pos_noise * rand_delta          # hand_default_dof_pos is the zero vector (:1245)
    dof = st["dof_state"].view(N, -1, 2)
    dof[ids, :D, 0] = pos
    dof[ids, :D, 1] = torch.zeros(D) + c.reset_dof_vel_noise * rf[:, 5 + D:5 + 2 * D]
    st["prev_targets"][ids, :D] = pos
    st["cur_targets"][ids, :D] = pos
    st["prev_prev_targets"][ids, :D] = pos
    # wrist-relative latches: the object pose as last observed (:841 reads the stale gather of compute_observations)
    palm = st["rigid_body_state"].view(N, -1, 13)[:, c.palm_body, 0:7]
    st["object_pose_wrt_wrist"] = pose_wrt_wrist(st["object_pose"], palm)
    st["obs_object_pose"][ids] = st["object_pose_wrt_wrist"][ids].clone()
    st["obs_object_pose_freq"][ids] = st["object_pose_wrt_wrist"][ids].clone()
    if len(ids) == N:
        st["progress_buf"] = torch.clamp((rnd[:, RND_PROGRESS] * c.max_episode_length).long(), max=c.max_episode_length - 1)
    else:
        st["progress_buf"][ids] = 0
    st["reset_buf"][ids] = 0
    st["apply_reset_buf"][ids] = 0
    st["successes"][ids] = 0.0
    st["best_rotation_dist"][ids] = -1.0
    st["hold_count_buf"][ids] = 0


def _apply_actions(st, c: DexConsts, actions: Tensor, rnd: Tensor, rna_perturbation: Optional[Tensor]) -> None:
    """:955-1015 with the ADR apply_action_noise_latency (:1430-1441)."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    lo, hi = st["dof_lower"], st["dof_upper"]
    st["actions"] = actions.clone()
    dof_pos = st["dof_state"].view(N, -1, 2)[:, :D, 0]
    refreshed = st["progress_buf"] == 0
    queue = st["prev_actions_queue"]
    queue[refreshed] = unscale(dof_pos[refreshed], lo, hi).view(-1, 1, D)
    st["prev_actions"][refreshed] = unscale(dof_pos[refreshed], lo, hi).view(-1, D)
    queue[:, 1:] = queue[:, :-1].clone()
    queue[:, 0, :] = st["actions"]
    delay = (rnd[:, RND_ACTION_DELAY] < st["adr_action_delay_prob"]).view(-1, 1)
    a = queue[torch.arange(N), st["action_latency"]] * ~delay + st["prev_actions"] * delay
    white = rnd[:, RND_ACTION_WHITE:RND_ACTION_WHITE + D] * gaussian_stdev(st["adr_affine_action_white"])
    a = st["affine_actions_scaling"] * a + st["affine_actions_additive"] + white
    if rna_perturbation is not None:                                   # :887-890 (the network itself is outside the path)
        alpha = st["adr_rna_alpha"].view(-1, 1)
        a = alpha * unscale(rna_perturbation, lo, hi) + (1 - alpha) * a
    prev, cur = st["prev_targets"], st["cur_targets"]
    if c.use_capped_dof_control:
        delta = scale(a, lo, hi) - prev
        cap = c.max_dof_radians_per_second * c.dt * c.control_freq_inv
        cur[:, :D] = prev + torch.clamp(delta, -cap, cap)
    else:
        cur[:, :D] = scale(a, lo, hi)
    cur[:, :D] = c.act_moving_average * cur + (1.0 - c.act_moving_average) * prev
    cur[:, :D] = torch.max(torch.min(cur, hi), lo)
    st["dof_delta"] = cur - prev
    st["prev_actions"][:] = st["actions"].clone()


def pre_step(st: Dict[str, Tensor], c: DexConsts, actions: Tensor, rnd: Tensor,
             rna_perturbation: Optional[Tensor] = None) -> Dict[str, Tensor]:
    """:912-949.  Mutates ``st``; returns the actor-row lists handed to the simulator.  ``apply_randomizations`` (host ADR) is
    outside the restatement: the envs in ``reset_buf & ~apply_reset_buf`` are only marked, their reset is applied one step later."""
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    mask_randomize = (st["reset_buf"] & ~st["apply_reset_buf"]).bool()
    ids = st["apply_reset_buf"].nonzero(as_tuple=False).flatten()
    randomized = mask_randomize.nonzero(as_tuple=False).flatten()
    if len(randomized) > 0:
        st["apply_reset_buf"][mask_randomize] = 1
    if len(goal_ids) > 0:
        _reset_target_pose(st, c, goal_ids, rnd[goal_ids, RND_GOAL:RND_GOAL + 7])
    root_rows = st["goal_object_indices"][goal_ids].to(torch.int32) if len(ids) == 0 else torch.zeros(0, dtype=torch.int32)
    hand_rows = torch.zeros(0, dtype=torch.int32)
    if len(ids) > 0:
        _reset_idx(st, c, ids, rnd)
        root_rows = torch.unique(torch.cat([st["object_indices"][ids], st["goal_object_indices"][ids],
                                            st["goal_object_indices"][goal_ids]]).to(torch.int32))
        hand_rows = st["hand_indices"][ids].to(torch.int32)
    _apply_actions(st, c, actions, rnd, rna_perturbation)
    if c.force_scale > 0.0:                                            # :1017-1030
        st["rb_forces"] *= torch.pow(torch.tensor(c.force_decay, dtype=torch.float32), c.dt / c.force_decay_interval)
        hit = (rnd[:, RND_FORCE_TEST] < st["random_force_prob"]).nonzero(as_tuple=False).squeeze(-1)
        st["rb_forces"][hit, c.object_body, :] = rnd[hit, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] * c.object_mass * c.force_scale
    return {"root_rows": root_rows, "hand_rows": hand_rows, "randomized": randomized}


# ---------------------------------------------------------------------------------------------
# post-physics step
# ---------------------------------------------------------------------------------------------
def _random_cube_observation(st, c: DexConsts, pose: Tensor, prnd: Tensor) -> Tensor:
    """:739-771 (random cube pose injection)."""
    rf = 2.0 * prnd[:, PRND_CUBE:PRND_CUBE + 5] + -1.0
    rot = _new_rot(c, rf[:, 3], rf[:, 4], prnd[:, PRND_CUBE_QUAT:PRND_CUBE_QUAT + 3])
    rc = st["random_cube_poses"]
    rc[:, 0:2] = st["object_init_state"][:, 0:2] + 0.5 * rf[:, 0:2]
    rc[:, 2] = st["object_init_state"][:, 2] + 0.5 * rf[:, 2]
    rc[:, 3:7] = rot
    mask = prnd[:, PRND_CUBE_MASK:PRND_CUBE_MASK + 1] < c.random_cube_pose_prob
    return pose * ~mask + rc * mask


def observe(st: Dict[str, Tensor], c: DexConsts, prnd: Tensor) -> None:
    """compute_observations, base (:606-692) then the ADR extension (:1443-1489).  Fills ``st["obs"]`` (unclamped)."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    lo, hi = st["dof_lower"], st["dof_upper"]
    obj = st["root_state"][st["object_indices"]]
    st["object_pose"] = obj[:, 0:7]
    goal_pose = st["goal_states"][:, 0:7]
    rb = st["rigid_body_state"].view(N, -1, 13)
    palm = rb[:, c.palm_body, 0:7]
    st["object_pose_wrt_wrist"] = pose_wrt_wrist(st["object_pose"], palm)
    goal_wrt_wrist = pose_wrt_wrist(goal_pose, palm)
    tips = rb[:, list(c.tip_bodies)][:, :, 0:13]
    st["frame"] = st["frame"] + 1
    dof = st["dof_state"].view(N, -1, 2)[:, :D]
    dof_pos, dof_vel = dof[..., 0], dof[..., 1]
    o: Dict[str, Tensor] = {}
    o["dof_pos"] = unscale(dof_pos, lo, hi)
    o["dof_vel"] = dof_vel.clone()
    o["dof_force"] = c.force_torque_obs_scale * st["dof_force"]
    o["object_pose"] = st["object_pose_wrt_wrist"].clone()
    o["object_vels"] = torch.cat([obj[:, 7:10], c.vel_obs_scale * obj[:, 10:13]], dim=-1)
    o["goal_pose"] = goal_wrt_wrist.clone()
    o["goal_relative_rot"] = quat_mul(st["object_pose_wrt_wrist"][:, 3:7], quat_conjugate(goal_wrt_wrist[:, 3:7]))
    o["ft_states"] = tips.reshape(N, 13 * c.n_tips)
    o["ft_force_torques"] = c.force_torque_obs_scale * st["sensors"]
    o["rb_forces"] = st["rb_forces"][:, [c.object_body], :].view(-1, 3)
    o["last_actions"] = st["actions"].clone()
    o["cube_random_params"] = st["cube_random_params"].clone()
    o["hand_random_params"] = st["hand_random_params"].clone()
    o["gravity_vec"] = st["gravity_vec"].clone()
    quat_diff = quat_mul(obj[:, 3:7], quat_conjugate(st["goal_states"][:, 3:7]))
    st["curr_rotation_dist"] = 2.0 * torch.asin(torch.clamp(torch.norm(quat_diff[:, 0:3], p=2, dim=-1), max=1.0))
    st["best_rotation_dist"] = torch.where(st["best_rotation_dist"] < 0.0, st["curr_rotation_dist"], st["best_rotation_dist"])
    o["rot_dist"] = torch.stack([st["curr_rotation_dist"], st["best_rotation_dist"]], dim=1)
    # ---- ADR extension ----
    update_freq = torch.remainder(st["frame"] + st["cube_pose_refresh_offset"], st["cube_pose_refresh_rate"]) == 0
    white_pose = prnd[:, PRND_POSE_WHITE:PRND_POSE_WHITE + 7] * gaussian_stdev(st["adr_affine_cube_pose_white"])
    noisy = _random_cube_observation(
        st, c, st["affine_cube_pose_scaling"] * st["object_pose_wrt_wrist"] + st["affine_cube_pose_additive"] + white_pose, prnd)
    st["obs_object_pose_freq"][update_freq] = noisy[update_freq]
    delay = prnd[:, PRND_DELAY] < st["adr_cube_obs_delay_prob"]
    st["obs_object_pose"][~delay] = st["obs_object_pose_freq"][~delay]
    white_dof = prnd[:, PRND_DOF_WHITE:PRND_DOF_WHITE + D] * gaussian_stdev(st["adr_affine_dof_pos_white"])
    randomized = st["affine_dof_pos_scaling"] * dof_pos + st["affine_dof_pos_additive"] + white_dof
    o["dof_pos_randomized"] = unscale(randomized, lo, hi)
    o["object_pose_cam_randomized"] = st["obs_object_pose"].clone()
    o["goal_relative_rot_cam_randomized"] = quat_mul(st["obs_object_pose"][:, 3:7], quat_conjugate(goal_wrt_wrist[:, 3:7]))
    o["stochastic_delay_params"] = torch.stack([st["adr_cube_obs_delay_prob"], st["cube_pose_refresh_rate"].float() / 6.0,
                                                st["adr_action_delay_prob"], st["action_latency"].float() / 60.0], dim=1)
    o["affine_params"] = torch.cat([st["affine_actions_scaling"], st["affine_actions_additive"], st["affine_cube_pose_scaling"],
                                    st["affine_cube_pose_additive"], st["affine_dof_pos_scaling"], st["affine_dof_pos_additive"]], dim=-1)
    st["obs"] = o
    st["obs_rows"] = torch.cat([o[k] for k, _ in OBS_KEYS], dim=-1)


def compute_hand_reward(st: Dict[str, Tensor], c: DexConsts) -> None:
    """:1598-1663 with the bookkeeping of compute_reward (:438-460)."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    obj = st["root_state"][st["object_indices"]]
    dof_vel = st["dof_state"].view(N, -1, 2)[:, :D, 1]
    goal_dist = torch.norm(obj[:, 0:3] - st["goal_states"][:, 0:3], p=2, dim=-1)
    quat_diff = quat_mul(obj[:, 3:7], quat_conjugate(st["goal_states"][:, 3:7]))
    rot_dist = 2.0 * torch.asin(torch.clamp(torch.norm(quat_diff[:, 0:3], p=2, dim=-1), max=1.0))
    dist_rew = goal_dist * c.dist_reward_scale
    rot_rew = 1.0 / (torch.abs(rot_dist) + c.rot_eps) * c.rot_reward_scale
    action_penalty = c.action_penalty_scale * torch.sum(st["actions"] ** 2, dim=-1)
    action_delta_penalty = c.action_delta_penalty_scale * torch.sum((st["cur_targets"] - st["prev_targets"]) ** 2, dim=-1)
    velocity_penalty = -0.05 * torch.sum((dof_vel / (5.0 - 1.0)) ** 2, dim=-1)
    rgb = st["reset_goal_buf"]
    goal_reached = torch.where(torch.abs(rot_dist) <= c.success_tolerance, torch.ones_like(rgb), rgb)
    hold = torch.where(goal_reached != 0, st["hold_count_buf"] + 1, torch.zeros_like(goal_reached))
    goal_resets = torch.where(hold > c.num_success_hold_steps, torch.ones_like(rgb), rgb)
    successes = st["successes"] + goal_resets
    reach_goal_rew = (goal_resets == 1) * c.reach_goal_bonus
    fall_rew = (goal_dist >= c.fall_dist) * c.fall_penalty
    resets = torch.where(goal_dist >= c.fall_dist, torch.ones_like(st["reset_buf"]), st["reset_buf"])
    progress = st["progress_buf"]
    if c.max_consecutive_successes > 0:
        progress = torch.where(torch.abs(rot_dist) <= c.success_tolerance, torch.zeros_like(progress), progress)
        resets = torch.where(successes >= c.max_consecutive_successes, torch.ones_like(resets), resets)
    timed_out = progress >= c.max_episode_length - 1
    resets = torch.where(timed_out, torch.ones_like(resets), resets)
    timeout_rew = timed_out * 0.5 * c.fall_penalty
    reward = dist_rew + rot_rew + action_penalty + action_delta_penalty + velocity_penalty + reach_goal_rew + fall_rew + timeout_rew
    num_resets = torch.sum(resets)
    finished = torch.sum(successes * resets.float())
    cons = st["consecutive_successes"]
    av = torch.tensor(c.av_factor, dtype=torch.float32)
    st["consecutive_successes"] = torch.where(num_resets > 0, av * finished / num_resets + (1.0 - av) * cons, cons)
    st["rew_buf"], st["reset_buf"], st["reset_goal_buf"], st["progress_buf"] = reward, resets, goal_resets, progress
    st["hold_count_buf"], st["successes"] = hold, successes
    terms = [dist_rew, rot_rew, action_penalty, action_delta_penalty, velocity_penalty, reach_goal_rew, fall_rew, timeout_rew]
    st["episode_cumulative"] = torch.stack([t.to(torch.float32) for t in terms], dim=0)
    st["best_rotation_dist"] = torch.minimum(st["best_rotation_dist"], st["curr_rotation_dist"])


def post_step(st: Dict[str, Tensor], c: DexConsts, prnd: Tensor) -> None:
    """post_physics_step (:1033-1048) + the VecTaskDextreme.step tail (adr_vec_task.py:289-325)."""
    D = c.n_dofs
    st["progress_buf"] = st["progress_buf"] + 1
    observe(st, c, prnd)
    compute_hand_reward(st, c)
    st["prev_targets"][:, :D] = st["cur_targets"][:, :D]
    st["timeout_buf"] = (st["progress_buf"] >= c.max_episode_length - 1) & (st["reset_buf"] != 0)
    st["obs_out"] = torch.clamp(st["obs_rows"], -c.clip_obs, c.clip_obs)


# ---------------------------------------------------------------------------------------------
# per-env ADR resampling on reset (the tensor half of apply_randomizations)
# ---------------------------------------------------------------------------------------------
# Tensorised ADR parameters of cfg/task/AllegroHandDextremeADR.yaml in declaration order, with their index among ALL adr params
# (13 physics parameters patched into Isaac Gym's DR ranges come first) -- adr_modes counts over all of them.
ADR_TENSOR_KEYS = ["cube_obs_delay_prob", "cube_pose_refresh_rate", "action_delay_prob", "action_latency", "affine_action_scaling",
                   "affine_action_additive", "affine_action_white", "affine_cube_pose_scaling", "affine_cube_pose_additive",
                   "affine_cube_pose_white", "affine_dof_pos_scaling", "affine_dof_pos_additive", "affine_dof_pos_white", "rna_alpha"]
ADR_N_BUILTIN = 13
ADR_N_PARAMS = ADR_N_BUILTIN + len(ADR_TENSOR_KEYS)
# random table of the resampling [n_envs, ARND_STRIDE]
ARND_WORKER = 0        # worker type draw
ARND_MODE = 1          # stands in for torch.randint(0, 2 * n_params): floor(u * 2 * n_params)
ARND_PARAM = 2         # one uniform per tensorised parameter (16 slots reserved)
ARND_LATENCY = 18
ARND_RATE = 19
ARND_OFFSET = 20
ARND_AFFINE = 21       # 16 + 16 + 7 + 7 + 16 + 16 normals
ARND_STRIDE = 99
ADR_ROLLOUT, ADR_BOUNDARY = 0, 1


@dataclasses.dataclass
class AdrTable:
    """Current ADR ranges (host state of adr_update) of the tensorised parameters."""
    ranges: Dict[str, Tuple[float, float]]
    next_limits: Dict[str, Tuple[float, float]]
    boundary_fraction: float = 0.4
    extended_boundary_sample: bool = False


def default_adr_table(widen: float = 0.0) -> AdrTable:
    """init_range of the YAML (``widen`` > 0 opens every range a little so that all branches produce distinct values)."""
    init = {"cube_obs_delay_prob": (0.0, 0.05), "cube_pose_refresh_rate": (1.0, 1.0), "action_delay_prob": (0.0, 0.05),
            "action_latency": (0.0, 0.0), "affine_action_scaling": (0.0, 0.0), "affine_action_additive": (0.0, 0.04),
            "affine_action_white": (0.0, 0.04), "affine_cube_pose_scaling": (0.0, 0.0), "affine_cube_pose_additive": (0.0, 0.04),
            "affine_cube_pose_white": (0.0, 0.04), "affine_dof_pos_scaling": (0.0, 0.0), "affine_dof_pos_additive": (0.0, 0.04),
            "affine_dof_pos_white": (0.0, 0.04), "rna_alpha": (0.0, 0.0)}
    if widen > 0.0:
        init = {k: (lo, hi + widen * (10.0 if k in ("cube_pose_refresh_rate", "action_latency") else 1.0)) for k, (lo, hi) in init.items()}
    nxt = {k: (max(lo - 0.01, 0.0), hi + 0.01) for k, (lo, hi) in init.items()}
    return AdrTable(ranges=init, next_limits=nxt)


def adr_resample(st: Dict[str, Tensor], c: DexConsts, table: AdrTable, mask: Tensor, arnd: Tensor) -> None:
    """recycle_envs (adr_vec_task.py:740-758) + sample_adr_tensor of every tensorised parameter (:666-730, as the tail of adr_update
    :915-917 calls it) + the Dextreme-specific resampling of allegro_hand_dextreme.py:1392-1416, for the envs in ``mask``."""
    ids = mask.nonzero(as_tuple=False).squeeze(-1)
    if len(ids) == 0:
        return
    D = c.n_dofs
    # recycle_envs
    new_types = torch.zeros(len(ids), dtype=torch.long)
    r = arnd[ids, ARND_WORKER]
    new_types[r < table.boundary_fraction] = ADR_ROLLOUT
    new_types[r >= table.boundary_fraction] = ADR_BOUNDARY
    st["worker_types"][ids] = new_types
    st["adr_modes"][ids] = torch.clamp((arnd[ids, ARND_MODE] * (2 * ADR_N_PARAMS)).long(), max=2 * ADR_N_PARAMS - 1)
    # sample_adr_tensor
    sample_mask = torch.zeros(mask.shape[0], dtype=torch.bool)
    sample_mask[ids] = True
    for t, name in enumerate(ADR_TENSOR_KEYS):
        n = ADR_N_BUILTIN + t
        lo, hi = table.ranges[name]
        nlo, nhi = table.next_limits[name]
        low = (st["worker_types"] == ADR_BOUNDARY) & (st["adr_modes"] == 2 * n) & sample_mask
        high = (st["worker_types"] == ADR_BOUNDARY) & (st["adr_modes"] == 2 * n + 1) & sample_mask
        rollout = (~low) & (~high) & sample_mask
        result = torch.zeros(len(ids), dtype=torch.float32)
        result[rollout[ids]] = arnd[rollout, ARND_PARAM + t] * (hi - lo) + lo
        result[low[ids]] = nlo if table.extended_boundary_sample else lo
        result[high[ids]] = nhi if table.extended_boundary_sample else hi
        st["adr_" + name][ids] = result
    # allegro_hand_dextreme.py:1398-1416
    st["action_latency"][ids] = (st["adr_action_latency"][ids] + (-(arnd[ids, ARND_LATENCY] - 0.5))).round().long()
    st["cube_pose_refresh_rate"][ids] = (st["adr_cube_pose_refresh_rate"][ids] + (-(arnd[ids, ARND_RATE] - 0.5))).round().long()
    st["cube_pose_refresh_offset"][mask] = (arnd[ids, ARND_OFFSET] * st["cube_pose_refresh_rate"][ids].view(-1).float() - 0.5).round().long()
    o = ARND_AFFINE
    for key, adr, dim, one in (("affine_actions_scaling", "affine_action_scaling", D, 1.), ("affine_actions_additive", "affine_action_additive", D, 0.),
                               ("affine_cube_pose_scaling", "affine_cube_pose_scaling", 7, 1.), ("affine_cube_pose_additive", "affine_cube_pose_additive", 7, 0.),
                               ("affine_dof_pos_scaling", "affine_dof_pos_scaling", D, 1.), ("affine_dof_pos_additive", "affine_dof_pos_additive", D, 0.)):
        noise = arnd[ids, o:o + dim] * gaussian_stdev(st["adr_" + adr][ids])
        st[key][ids] = (1. + noise) if one else noise
        o += dim
