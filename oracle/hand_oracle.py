"""ORACLE (test infrastructure, not product code) -- CPU restatement of the ShadowHand / AllegroHand hot path.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU legs may import this.

Torch-on-CPU restatement, op for op in the reference's association order, of

* ``pre_step``  <- ``shadow_hand.py:670-708`` / ``allegro_hand.py:586-625`` (pre_physics_step),
                   ``shadow_hand.py:604-668`` / ``allegro_hand.py:524-584`` (reset_idx),
                   ``shadow_hand.py:586-602`` / ``allegro_hand.py:506-522`` (reset_target_pose),
                   ``shadow_hand.py:803-806`` (randomize_rotation), ``torch_jit_utils.py:118-123`` (quat_from_angle_axis)
* ``post_step`` <- ``shadow_hand.py:710-715`` / ``allegro_hand.py:627-632`` (post_physics_step),
                   ``shadow_hand.py:437-584`` / ``allegro_hand.py:406-504`` (observations, all four layouts + asymmetric states),
                   ``shadow_hand.py:746-800`` / ``allegro_hand.py:663-719`` (compute_hand_reward), ``vec_task.py:424-432`` (step tail)

Parity pinning: ``tests/test_hand_oracle_vs_reference.py`` runs the reference's real ``ShadowHand`` /
``AllegroHand`` classes (through ``oracle/reference_loader.py``) and this restatement on the same inputs
whenever ``/root/reference`` is present; ``tests/golden/hand_*.npz`` hold reference-generated vectors
(``oracle/make_golden.py``).  Random numbers come from a per-env table laid out as ``HA_HAND_RND_*``.
"""
from __future__ import annotations

import dataclasses
from typing import Dict, Tuple

import numpy as np
import torch

from .kuka_oracle import scale, unscale

Tensor = torch.Tensor

# per-env random table (include/handarm_b200.h, HA_HAND_RND_*): raw U(0,1) draws, N(0,1) in the force-normal slots
RND_GOAL, RND_RESET_GOAL, RND_FORCE_PROB, RND_FORCE_TEST, RND_FORCE_NORMAL, RND_RESET = 0, 4, 8, 9, 10, 13

OBS_OPENAI, OBS_FULL_NO_VEL, OBS_FULL, OBS_FULL_STATE = 0, 1, 2, 3
OBS_NAMES = {"openai": OBS_OPENAI, "full_no_vel": OBS_FULL_NO_VEL, "full": OBS_FULL, "full_state": OBS_FULL_STATE}


def rnd_stride(n_dofs: int) -> int:
    return 13 + 2 * n_dofs + 5


@dataclasses.dataclass
class HandConsts:
    """Host scalars of a hand task (python floats, the reference's attributes). Defaults: ShadowHand.yaml."""
    kind: str = "shadow"                  # "shadow" | "allegro"
    n_dofs: int = 24
    n_actions: int = 20
    actuated: Tuple[int, ...] = (0, 1, 2, 3, 4, 6, 7, 8, 10, 11, 12, 14, 15, 16, 17, 19, 20, 21, 22, 23)
    n_tips: int = 5
    tip_bodies: Tuple[int, ...] = (7, 11, 15, 20, 25)
    n_bodies: int = 28
    object_body: int = 26
    obs_type: str = "full_state"
    asymmetric: bool = False
    dt: float = 0.01667
    max_episode_length: int = 600
    dist_reward_scale: float = -10.0
    rot_reward_scale: float = 1.0
    rot_eps: float = 0.1
    action_penalty_scale: float = -0.0002
    success_tolerance: float = 0.1
    reach_goal_bonus: float = 250.0
    fall_dist: float = 0.24
    fall_penalty: float = 0.0
    max_consecutive_successes: int = 0
    av_factor: float = 0.1
    ignore_z_rot: bool = False
    vel_obs_scale: float = 0.2
    force_torque_obs_scale: float = 10.0
    use_relative_control: bool = False
    dof_speed_scale: float = 20.0
    act_moving_average: float = 1.0
    force_scale: float = 0.0
    force_decay: float = 0.99
    force_decay_interval: float = 0.08
    force_prob_range: Tuple[float, float] = (0.001, 0.1)
    object_mass: float = 0.1
    reset_position_noise: float = 0.01
    reset_dof_pos_noise: float = 0.2
    reset_dof_vel_noise: float = 0.0
    goal_displacement: Tuple[float, float, float] = (-0.2, -0.06, 0.12)
    clip_obs: float = 5.0
    up_axis_idx: int = 2

    @property
    def obs_dim(self) -> int:
        return hand_obs_dim(self.kind, self.obs_type)

    @property
    def states_dim(self) -> int:
        return hand_obs_dim(self.kind, "full_state") if self.asymmetric else 0


def allegro_consts(**kw) -> HandConsts:
    """AllegroHand.yaml / allegro_hand.py: 16 DOFs all actuated, no fingertip observations."""
    base = dict(kind="allegro", n_dofs=16, n_actions=16, actuated=tuple(range(16)), n_tips=0, tip_bodies=(), n_bodies=19, object_body=17)
    base.update(kw)
    return HandConsts(**base)


def hand_obs_dim(kind: str, obs_type: str) -> int:
    """shadow_hand.py:111-116, allegro_hand.py:108-112."""
    if kind == "shadow":
        return {"openai": 42, "full_no_vel": 77, "full": 157, "full_state": 211}[obs_type]
    return {"full_no_vel": 50, "full": 72, "full_state": 88}[obs_type]


# ---------------------------------------------------------------------------------------------
# math helpers (torch_jit_utils.py)
# ---------------------------------------------------------------------------------------------
def quat_mul(a: Tensor, b: Tensor) -> Tensor:
    """torch_jit_utils.py:41-62."""
    x1, y1, z1, w1 = a[:, 0], a[:, 1], a[:, 2], a[:, 3]
    x2, y2, z2, w2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    ww = (z1 + x1) * (x2 + y2)
    yy = (w1 - y1) * (w2 + z2)
    zz = (w1 + y1) * (w2 - z2)
    xx = ww + yy + zz
    qq = 0.5 * (xx + (z1 - x1) * (x2 - y2))
    w = qq - ww + (z1 - y1) * (y2 - z2)
    x = qq - xx + (x1 + w1) * (x2 + w2)
    y = qq - yy + (w1 - x1) * (y2 + z2)
    z = qq - zz + (z1 + y1) * (w2 - x2)
    return torch.stack([x, y, z, w], dim=-1)


def quat_conjugate(a: Tensor) -> Tensor:
    return torch.cat((-a[:, :3], a[:, -1:]), dim=-1)


def _normalize(x: Tensor) -> Tensor:
    return x / x.norm(p=2, dim=-1).clamp(min=1e-9, max=None).unsqueeze(-1)


def quat_from_angle_axis(angle: Tensor, axis: Tensor) -> Tensor:
    """torch_jit_utils.py:118-123."""
    theta = (angle / 2).unsqueeze(-1)
    xyz = _normalize(axis) * theta.sin()
    w = theta.cos()
    return _normalize(torch.cat([xyz, w], dim=-1))


def randomize_rotation(rand0: Tensor, rand1: Tensor) -> Tensor:
    """shadow_hand.py:803-806 with the x / y unit tensors."""
    n = rand0.shape[0]
    x_unit = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    y_unit = torch.tensor([0.0, 1.0, 0.0]).repeat(n, 1)
    return quat_mul(quat_from_angle_axis(rand0 * np.pi, x_unit), quat_from_angle_axis(rand1 * np.pi, y_unit))


def randomize_rotation_pen(rand0: Tensor, rand1: Tensor) -> Tensor:
    """shadow_hand.py:809-814 as reset_idx calls it (:626-629, max_angle = tensor(0.3)); rand1 is unused by the reference too."""
    n = rand0.shape[0]
    x_unit = torch.tensor([1.0, 0.0, 0.0]).repeat(n, 1)
    z_unit = torch.tensor([0.0, 0.0, 1.0]).repeat(n, 1)
    max_angle = torch.tensor(0.3)
    return quat_mul(quat_from_angle_axis(0.5 * np.pi + rand0 * max_angle, x_unit), quat_from_angle_axis(rand0 * np.pi, z_unit))


# ---------------------------------------------------------------------------------------------
# pre-physics step
# ---------------------------------------------------------------------------------------------
def _reset_target_pose(st: Dict[str, Tensor], c: HandConsts, ids: Tensor, u4: Tensor) -> None:
    """shadow_hand.py:586-602 / allegro_hand.py:506-522; ``u4`` raw U(0,1) draws [n, 4]."""
    r = 2.0 * u4 + -1.0                                   # torch_rand_float(-1, 1): (hi - lo) * rand + lo
    new_rot = randomize_rotation(r[:, 0], r[:, 1])
    st["goal_states"][ids, 0:3] = st["goal_init_state"][ids, 0:3]
    st["goal_states"][ids, 3:7] = new_rot
    rows = st["goal_object_indices"][ids]
    st["root_state"][rows, 0:3] = st["goal_states"][ids, 0:3] + torch.tensor(c.goal_displacement, dtype=torch.float32)
    st["root_state"][rows, 3:7] = st["goal_states"][ids, 3:7]
    st["root_state"][rows, 7:13] = 0.0
    st["reset_goal_buf"][ids] = 0


def _reset_idx(st: Dict[str, Tensor], c: HandConsts, ids: Tensor, rnd: Tensor) -> None:
    """shadow_hand.py:604-668 / allegro_hand.py:524-584."""
    D = c.n_dofs
    r = 2.0 * rnd[ids, RND_RESET:RND_RESET + 2 * D + 5] + -1.0
    _reset_target_pose(st, c, ids, rnd[ids, RND_RESET_GOAL:RND_RESET_GOAL + 4])
    st["rb_forces"][ids] = 0.0
    rows = st["object_indices"][ids]
    init = st["object_init_state"][ids]
    st["root_state"][rows] = init.clone()
    st["root_state"][rows, 0:2] = init[:, 0:2] + c.reset_position_noise * r[:, 0:2]
    st["root_state"][rows, c.up_axis_idx] = init[:, c.up_axis_idx] + c.reset_position_noise * r[:, c.up_axis_idx]
    # objectType == "pen" (ignore_z_rot): the pen starts roughly upright (shadow_hand.py:626-629)
    st["root_state"][rows, 3:7] = randomize_rotation_pen(r[:, 3], r[:, 4]) if c.ignore_z_rot else randomize_rotation(r[:, 3], r[:, 4])
    st["root_state"][rows, 7:13] = 0.0
    pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
    st["random_force_prob"][ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, RND_FORCE_PROB] + torch.log(pr[1]))
    lo, hi, dflt = st["dof_lower"], st["dof_upper"], st["dof_default"]
    delta_max = hi - dflt
    delta_min = lo - dflt
    rand_delta = delta_min + (delta_max - delta_min) * 0.5 * (r[:, 5:5 + D] + 1)
    pos = dflt + c.reset_dof_pos_noise * rand_delta
    st["dof_state"][ids, :D, 0] = pos
    st["dof_state"][ids, :D, 1] = torch.zeros(D) + c.reset_dof_vel_noise * r[:, 5 + D:5 + 2 * D]
    st["prev_targets"][ids, :D] = pos
    st["cur_targets"][ids, :D] = pos
    st["progress_buf"][ids] = 0
    st["reset_buf"][ids] = 0
    st["successes"][ids] = 0


def pre_step(st: Dict[str, Tensor], c: HandConsts, actions: Tensor, rnd: Tensor) -> Dict[str, Tensor]:
    """shadow_hand.py:670-708 / allegro_hand.py:586-625.  Mutates ``st``; returns the actor-row lists the
    reference hands to ``set_actor_root_state_tensor_indexed`` / ``set_dof_*_indexed``."""
    ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    root_rows = torch.zeros(0, dtype=torch.int32)
    hand_rows = torch.zeros(0, dtype=torch.int32)
    if len(goal_ids) > 0:
        _reset_target_pose(st, c, goal_ids, rnd[goal_ids, RND_GOAL:RND_GOAL + 4])
        if len(ids) == 0:
            root_rows = st["goal_object_indices"][goal_ids].to(torch.int32)
    if len(ids) > 0:
        _reset_idx(st, c, ids, rnd)
        root_rows = torch.unique(torch.cat([st["object_indices"][ids], st["goal_object_indices"][ids],
                                            st["goal_object_indices"][goal_ids]]).to(torch.int32))
        hand_rows = st["hand_indices"][ids].to(torch.int32)
    st["actions"] = actions.clone()
    act = torch.tensor(c.actuated, dtype=torch.long)
    lo, hi = st["dof_lower"][act], st["dof_upper"][act]
    prev, cur = st["prev_targets"], st["cur_targets"]
    if c.use_relative_control:
        targets = prev[:, act] + c.dof_speed_scale * c.dt * actions
        cur[:, act] = torch.max(torch.min(targets, hi), lo)
    else:
        cur[:, act] = scale(actions, lo, hi)
        cur[:, act] = c.act_moving_average * cur[:, act] + (1.0 - c.act_moving_average) * prev[:, act]
        cur[:, act] = torch.max(torch.min(cur[:, act], hi), lo)
    prev[:, act] = cur[:, act]
    if c.force_scale > 0.0:
        st["rb_forces"] *= torch.pow(torch.tensor(c.force_decay, dtype=torch.float32), c.dt / c.force_decay_interval)
        hit = (rnd[:, RND_FORCE_TEST] < st["random_force_prob"]).nonzero(as_tuple=False).squeeze(-1)
        st["rb_forces"][hit, c.object_body, :] = rnd[hit, RND_FORCE_NORMAL:RND_FORCE_NORMAL + 3] * c.object_mass * c.force_scale
    return {"root_rows": torch.sort(root_rows).values, "hand_rows": hand_rows}


# ---------------------------------------------------------------------------------------------
# post-physics step
# ---------------------------------------------------------------------------------------------
def _full_state_row(st, c: HandConsts, q) -> Tensor:
    """compute_full_state (shadow_hand.py:528-584 / allegro_hand.py:462-504) as one [N, W] tensor."""
    D = c.n_dofs
    parts = [unscale(q["dof_pos"], st["dof_lower"], st["dof_upper"]), c.vel_obs_scale * q["dof_vel"],
             c.force_torque_obs_scale * st["dof_force"], q["object_pose"], q["object_linvel"], c.vel_obs_scale * q["object_angvel"],
             q["goal_pose"], q["quat_diff"]]
    if c.kind == "shadow":
        parts += [q["fingertip_state"].reshape(-1, 13 * c.n_tips), c.force_torque_obs_scale * st["sensors"]]
    parts.append(st["actions"])
    return torch.cat(parts, dim=-1)


def observations(st: Dict[str, Tensor], c: HandConsts) -> None:
    """compute_observations: fills ``obs_buf`` (and ``states_buf`` when asymmetric)."""
    N, D = st["reset_buf"].shape[0], c.n_dofs
    obj = st["root_state"][st["object_indices"]]
    dof = st["dof_state"].view(N, -1, 2)[:, :D]
    q = dict(dof_pos=dof[..., 0], dof_vel=dof[..., 1], object_pose=obj[:, 0:7], object_linvel=obj[:, 7:10], object_angvel=obj[:, 10:13],
             goal_pose=st["goal_states"][:, 0:7])
    q["quat_diff"] = quat_mul(obj[:, 3:7], quat_conjugate(st["goal_states"][:, 3:7]))
    if c.kind == "shadow":
        rb = st["rigid_body_state"].view(N, -1, 13)
        q["fingertip_state"] = rb[:, list(c.tip_bodies)][:, :, 0:13]
        q["fingertip_pos"] = rb[:, list(c.tip_bodies)][:, :, 0:3]
    un = unscale(q["dof_pos"], st["dof_lower"], st["dof_upper"])
    t = c.obs_type
    if t == "openai":                                   # shadow_hand.py:473-496 (no_vel=True)
        row = torch.cat([q["fingertip_pos"].reshape(N, 15), q["object_pose"][:, 0:3], q["quat_diff"], st["actions"]], dim=-1)
    elif t == "full_no_vel":                            # shadow_hand.py:499-511 / allegro_hand.py:438-447
        parts = [un, q["object_pose"], q["goal_pose"], q["quat_diff"]]
        if c.kind == "shadow":
            parts.append(q["fingertip_pos"].reshape(N, 15))
        row = torch.cat(parts + [st["actions"]], dim=-1)
    elif t == "full":                                   # shadow_hand.py:512-526 / allegro_hand.py:448-460
        parts = [un, c.vel_obs_scale * q["dof_vel"], q["object_pose"], q["object_linvel"], c.vel_obs_scale * q["object_angvel"],
                 q["goal_pose"], q["quat_diff"]]
        if c.kind == "shadow":
            parts.append(q["fingertip_state"].reshape(N, 65))
        row = torch.cat(parts + [st["actions"]], dim=-1)
    else:
        row = _full_state_row(st, c, q)
    assert row.shape[1] == c.obs_dim, (row.shape, c.obs_dim)
    st["obs_buf"] = row.contiguous()
    if c.asymmetric:
        st["states_buf"] = _full_state_row(st, c, q).contiguous()


def compute_hand_reward(st: Dict[str, Tensor], c: HandConsts) -> None:
    """shadow_hand.py:746-800 / allegro_hand.py:663-719."""
    obj = st["root_state"][st["object_indices"]]
    object_pos, object_rot = obj[:, 0:3], obj[:, 3:7]
    target_pos, target_rot = st["goal_states"][:, 0:3], st["goal_states"][:, 3:7]
    actions = st["actions"]
    reset_buf, reset_goal_buf, progress_buf = st["reset_buf"], st["reset_goal_buf"], st["progress_buf"]
    successes, consecutive_successes = st["successes"], st["consecutive_successes"]
    goal_dist = torch.norm(object_pos - target_pos, p=2, dim=-1)
    success_tolerance = c.success_tolerance
    if c.ignore_z_rot:
        success_tolerance = 2.0 * success_tolerance
    quat_diff = quat_mul(object_rot, quat_conjugate(target_rot))
    rot_dist = 2.0 * torch.asin(torch.clamp(torch.norm(quat_diff[:, 0:3], p=2, dim=-1), max=1.0))
    dist_rew = goal_dist * c.dist_reward_scale
    rot_rew = 1.0 / (torch.abs(rot_dist) + c.rot_eps) * c.rot_reward_scale
    action_penalty = torch.sum(actions ** 2, dim=-1)
    reward = dist_rew + rot_rew + action_penalty * c.action_penalty_scale
    goal_resets = torch.where(torch.abs(rot_dist) <= success_tolerance, torch.ones_like(reset_goal_buf), reset_goal_buf)
    successes = successes + goal_resets
    reward = torch.where(goal_resets == 1, reward + c.reach_goal_bonus, reward)
    reward = torch.where(goal_dist >= c.fall_dist, reward + c.fall_penalty, reward)
    resets = torch.where(goal_dist >= c.fall_dist, torch.ones_like(reset_buf), reset_buf)
    if c.max_consecutive_successes > 0:
        progress_buf = torch.where(torch.abs(rot_dist) <= success_tolerance, torch.zeros_like(progress_buf), progress_buf)
        resets = torch.where(successes >= c.max_consecutive_successes, torch.ones_like(resets), resets)
    timed_out = progress_buf >= float(c.max_episode_length) - 1
    resets = torch.where(timed_out, torch.ones_like(resets), resets)
    if c.max_consecutive_successes > 0:
        reward = torch.where(timed_out, reward + 0.5 * c.fall_penalty, reward)
    num_resets = torch.sum(resets)
    finished_cons_successes = torch.sum(successes * resets.float())
    av = torch.tensor(c.av_factor, dtype=torch.float32)
    cons = torch.where(num_resets > 0, av * finished_cons_successes / num_resets + (1.0 - av) * consecutive_successes, consecutive_successes)
    st["rew_buf"], st["reset_buf"], st["reset_goal_buf"], st["progress_buf"] = reward, resets, goal_resets, progress_buf
    st["successes"], st["consecutive_successes"] = successes, cons


def post_step(st: Dict[str, Tensor], c: HandConsts) -> None:
    """post_physics_step + the VecTask.step tail (vec_task.py:424-432)."""
    st["progress_buf"] = st["progress_buf"] + 1
    st["randomize_buf"] = st["randomize_buf"] + 1
    observations(st, c)
    compute_hand_reward(st, c)
    st["timeout_buf"] = (st["progress_buf"] >= c.max_episode_length - 1) & (st["reset_buf"] != 0)
    # what step() hands to the caller: clamp(obs_buf, +-clipObservations) as a new tensor (vec_task.py:432, :374-376)
    st["obs_out"] = torch.clamp(st["obs_buf"], -c.clip_obs, c.clip_obs)
    if c.asymmetric:
        st["states_out"] = torch.clamp(st["states_buf"], -c.clip_obs, c.clip_obs)
