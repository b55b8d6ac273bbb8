"""AllegroKuka single-arm tasks on the fused sm_100a kernels.

Drop-in for ``isaacgymenvs/tasks/allegro_kuka/allegro_kuka_base.py`` (``AllegroKukaBase``) and
``allegro_kuka_reorientation.py`` / ``_regrasping.py`` / ``_throw.py``: same constructor, same
attribute names for every tensor a caller can see, same ``extras`` keys.  ``pre_physics_step`` and
``post_physics_step`` are one kernel launch each (``ha_kuka_pre_step`` / ``ha_kuka_post_step``);
nothing on the step path synchronises with the host.
"""
from __future__ import annotations

import ctypes
import math
from typing import Any, Dict, List, Optional

import numpy as np
import torch

from .. import _lib
from ..sim import (KUKA2_LAYOUT, KUKA_ALLEGRO_LOWER, KUKA_ALLEGRO_UPPER, KUKA_LAYOUT, ActorLayout, SyntheticSim, cuboid_scales)
from ..vec_task import VecTask

C = _lib.CONSTS

EPISODE_TERM_KEYS = [
    "raw_fingertip_delta_rew", "raw_hand_delta_penalty", "raw_lifting_rew", "raw_keypoint_rew", "fingertip_delta_rew",
    "hand_delta_penalty", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "bonus_rew", "kuka_actions_penalty",
    "allegro_actions_penalty",
]   # allegro_kuka_base.py:355-368
STEP_TERM_KEYS = [
    "fingertip_delta_rew", "hand_delta_penalty", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "kuka_actions_penalty",
    "allegro_actions_penalty", "bonus_rew",
]   # allegro_kuka_base.py:912-921
EPISODE_TERM_KEYS_2ARMS = ["raw_fingertip_delta_rew", "raw_lifting_rew", "raw_keypoint_rew", "fingertip_delta_rew", "lifting_rew",
                           "lift_bonus_rew", "keypoint_rew", "bonus_rew"]                      # allegro_kuka_two_arms.py:345-354
STEP_TERM_KEYS_2ARMS = ["fingertip_delta_rew", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "bonus_rew"]   # :861-867


def tolerance_curriculum(last_update: int, frames: int, interval: int, prev_episode_successes: torch.Tensor, tol: float,
                         initial_tol: float, target_tol: float, increment: float):
    """Host-side curriculum, same rule as ``allegro_kuka_utils.py:86-115`` (one reduction every ``interval`` frames)."""
    if frames - last_update < interval:
        return tol, last_update
    if float(prev_episode_successes.mean()) < 3.0:
        return tol, last_update
    tol = max(min(tol * increment, initial_tol), target_tol)
    return tol, frames


def true_objective_coeffs(tol: float, initial_tol: float, target_tol: float):
    """``true_objective = successes * scale + offset`` (``allegro_kuka_utils.py:128-158``)."""
    offset = (initial_tol - tol) / (initial_tol - target_tol) if initial_tol > target_tol else 1.0
    scale = 0.01 if tol > target_tol else 1.0
    return scale, offset


class AllegroKukaBase(VecTask):
    subtask_id = C["HA_KUKA_REORIENTATION"]
    num_arms = 1                      # AllegroKukaTwoArmsBase (allegro_kuka_two_arms.py) sets 2
    default_layout = KUKA_LAYOUT
    episode_term_keys, step_term_keys = EPISODE_TERM_KEYS, STEP_TERM_KEYS

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture=False, force_render=False,
                 sim: Optional[SyntheticSim] = None, layout: Optional[ActorLayout] = None, seed: int = 1234, snapshots: int = 1):
        self.cfg = cfg
        env = cfg["env"]
        self._injected_sim, self._layout, self._seed, self._snapshots = sim, layout or self.default_layout, seed, snapshots
        self.frame_since_restart = 0
        self.clamp_abs_observations = env["clampAbsObservations"]
        self.privileged_actions = env["privilegedActions"]
        if self.privileged_actions:
            raise NotImplementedError("privilegedActions is off in every shipped config and not built here")
        self.num_arm_dofs, self.num_finger_dofs, self.num_allegro_fingertips = 7, 4, 4
        self.num_hand_dofs = self.num_finger_dofs * self.num_allegro_fingertips
        self.num_hand_arm_dofs = self.num_hand_dofs + self.num_arm_dofs
        self.num_allegro_kuka_actions = self.num_hand_arm_dofs * self.num_arms
        self.randomize = cfg["task"]["randomize"]
        self.randomization_params = cfg["task"]["randomization_params"]
        self.distance_delta_rew_scale = env["distanceDeltaRewScale"]
        self.lifting_rew_scale = env["liftingRewScale"]
        self.lifting_bonus = env["liftingBonus"]
        self.lifting_bonus_threshold = env["liftingBonusThreshold"]
        self.keypoint_rew_scale = env["keypointRewScale"]
        self.kuka_actions_penalty_scale = env["kukaActionsPenaltyScale"]
        self.allegro_actions_penalty_scale = env["allegroActionsPenaltyScale"]
        self.initial_tolerance = env["successTolerance"]
        self.success_tolerance = self.initial_tolerance
        self.target_tolerance = env["targetSuccessTolerance"]
        self.tolerance_curriculum_increment = env["toleranceCurriculumIncrement"]
        self.tolerance_curriculum_interval = env["toleranceCurriculumInterval"]
        self.reach_goal_bonus = env["reachGoalBonus"]
        self.fall_dist, self.fall_penalty = env["fallDistance"], env["fallPenalty"]
        self.reset_position_noise_x = env["resetPositionNoiseX"]
        self.reset_position_noise_y = env["resetPositionNoiseY"]
        self.reset_position_noise_z = env["resetPositionNoiseZ"]
        self.reset_rotation_noise = env["resetRotationNoise"]
        self.reset_dof_pos_noise_fingers = env["resetDofPosRandomIntervalFingers"]
        self.reset_dof_pos_noise_arm = env["resetDofPosRandomIntervalArm"]
        self.reset_dof_vel_noise = env["resetDofVelRandomInterval"]
        self.force_scale = env.get("forceScale", 0.0)
        self.force_prob_range = env.get("forceProbRange", [0.001, 0.1])
        self.force_decay = env.get("forceDecay", 0.99)
        self.force_decay_interval = env.get("forceDecayInterval", 0.08)
        self.hand_dof_speed_scale = env["dofSpeedScale"]
        self.use_relative_control = env["useRelativeControl"]
        if self.use_relative_control:
            raise NotImplementedError("Use relative control False for now")   # allegro_kuka_base.py:1373-1374
        self.act_moving_average = env["actionsMovingAverage"]
        self.debug_viz = False
        self.max_episode_length = env["episodeLength"]
        self.reset_time = env.get("resetTime", -1.0)
        self.max_consecutive_successes = env["maxConsecutiveSuccesses"]
        self.success_steps = int(env["successSteps"])
        self.keypoint_scale = env["keypointScale"]
        self.object_base_size = env["objectBaseSize"]
        self.randomize_object_dimensions = env["randomizeObjectDimensions"]
        self.with_small_cuboids, self.with_big_cuboids, self.with_sticks = env["withSmallCuboids"], env["withBigCuboids"], env["withSticks"]
        if self.reset_time > 0.0:
            self.max_episode_length = int(round(self.reset_time / (env.get("controlFrequencyInv", 1) * cfg["sim"]["dt"])))
        self.object_type = env["objectType"]
        assert self.object_type in ["block"]
        self.keypoints_offsets = self._object_keypoint_offsets()
        self.num_keypoints = len(self.keypoints_offsets)
        self.fingertip_offsets = np.array([[0.05, 0.005, 0], [0.05, 0.005, 0], [0.05, 0.005, 0], [0.06, 0.005, 0]], dtype=np.float32)
        self.palm_offset = np.array([-0.00, -0.02, 0.16], dtype=np.float32)
        self.obs_type = env["observationType"]
        if self.obs_type not in ["full_state"]:
            raise Exception("Unknown type of observations!")            # allegro_kuka_base.py:180-181
        T, K, D = self.num_allegro_fingertips, self.num_keypoints, self.num_hand_arm_dofs
        if self.num_arms == 1:
            self.full_state_size = 2 * D + 3 + 10 + 10 + 3 * T + 6 * K + 3 + 1 + 1 + 2 + T + 1   # allegro_kuka_base.py:185-220
        else:   # allegro_kuka_two_arms.py:176-211: no closest_fingertip_dist slots, keypoints relative to both palms
            A = self.num_arms
            self.full_state_size = 2 * D * A + 3 * A + 10 * A + 10 + 3 * T * A + 3 * K * A + 3 * K + 3 + 1 + 1 + 2 + 1
        self.num_obs_dict = {"full_state": self.full_state_size}
        self.up_axis, self.up_axis_idx = "z", 2
        cfg["env"]["numObservations"] = self.num_obs_dict[self.obs_type]
        cfg["env"]["numStates"] = self.full_state_size
        cfg["env"]["numActions"] = self.num_allegro_kuka_actions
        cfg["device_type"] = sim_device.split(":")[0]
        cfg["device_id"] = int(sim_device.split(":")[1]) if ":" in sim_device else 0
        cfg["headless"] = headless
        super().__init__(cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture, force_render)
        self._allocate_task_state()
        self._build_abi()
        # recorded initial states (allegro_kuka_base.py:98-101, :545-546; the two-arm base has no such option)
        self.should_load_initial_states = bool(env.get("loadInitialStates", False)) and self.num_arms == 1
        self.load_states_filename = env.get("loadStatesFile", "")
        self.initial_root_state_tensors = self.initial_dof_state_tensors = None
        self.initial_state_idx = self.num_initial_states = 0
        if self.should_load_initial_states:
            self.load_initial_states()

    # ---- subtask interface (allegro_kuka_base.py:396-470) ---------------------------------------
    def _object_keypoint_offsets(self):
        raise NotImplementedError()

    # ---- construction ------------------------------------------------------------------------
    def create_sim(self):
        """Replaces asset loading / actor creation (allegro_kuka_base.py:544-751) by the synthetic producer."""
        self.dt = self.sim_params_dt
        L = self._layout
        self.sim = self._injected_sim or SyntheticSim(self.num_envs, L, self.device, seed=self._seed, snapshots=self._snapshots)
        self.gym = self.sim
        N, dev = self.num_envs, self.device
        self.num_bodies = L.bodies_per_env
        self.num_dofs = L.dofs_per_env
        self.arm_hand_dof_lower_limits = torch.tensor(KUKA_ALLEGRO_LOWER * self.num_arms, dtype=torch.float, device=dev)
        self.arm_hand_dof_upper_limits = torch.tensor(KUKA_ALLEGRO_UPPER * self.num_arms, dtype=torch.float, device=dev)
        self.allegro_palm_handle = L.palm_body[0]
        self.allegro_palm_handles = list(L.palm_body)
        self.allegro_fingertip_handles = torch.tensor(list(L.tip_body), dtype=torch.long, device=dev)
        self.object_rb_handles = torch.tensor([L.object_body], dtype=torch.long, device=dev)
        self.object_rb_masses = torch.tensor([400.0 * self.object_base_size ** 3], dtype=torch.float, device=dev)
        ar = torch.arange(N, dtype=torch.long, device=dev) * L.actors_per_env
        self.allegro_hand_indices = ar + L.hand_actor[0]
        self.arm_indices = torch.stack([ar + h for h in L.hand_actor], dim=1)     # [N, num_arms] (allegro_kuka_two_arms.py:661-680)
        self.object_indices = ar + L.object_actor
        self.goal_object_indices = ar + L.goal_actor
        scales = cuboid_scales(self.with_small_cuboids, self.with_big_cuboids, self.with_sticks)
        if not self.randomize_object_dimensions:
            scales = [[1.0, 1.0, 1.0]]
        tab = torch.tensor(scales, dtype=torch.float)
        self.object_scales = tab[torch.arange(N) % len(scales)].contiguous().to(dev)
        kp = torch.tensor(self.keypoints_offsets, dtype=torch.float)                                   # [K,3]
        # keypoint[c] *= scale[c] * base_size * keypoint_scale / 2 (allegro_kuka_base.py:704-709), in python doubles
        ofs = np.asarray(self.keypoints_offsets, dtype=np.float64)[None] * (np.asarray(scales, dtype=np.float64)[:, None, :]
                                                                              * self.object_base_size * self.keypoint_scale / 2)
        tab_ofs = torch.tensor(ofs, dtype=torch.float)
        self.object_keypoint_offsets = tab_ofs[torch.arange(N) % len(scales)].contiguous().to(dev)   # [N,K,3]
        # object start pose: allegro (0, 0.8, 0) + table (dy -0.8, dz 0.38 + 0.25)  (allegro_kuka_base.py:402-412, :600-606)
        init = torch.zeros(N, 13, dtype=torch.float)
        init[:, 0:3] = torch.tensor([0.0, 0.8 - 0.8, 0.0 + 0.38 + 0.25])
        init[:, 6] = 1.0
        self.object_init_state = init.to(dev)
        self.goal_states = self.object_init_state.clone()
        self.goal_states[:, self.up_axis_idx] -= 0.04
        self.goal_init_state = self.goal_states.clone()

    def _allocate_task_state(self):
        """allegro_kuka_base.py:258-372, same names and dtypes."""
        N, dev, T = self.num_envs, self.device, self.num_allegro_fingertips
        f = dict(dtype=torch.float, device=dev)
        if self.num_arms == 1:
            self.target_volume_origin = torch.tensor([0, 0.05, 0.8], **f)
            self.target_volume_extent = torch.tensor([[-0.4, 0.4], [-0.05, 0.3], [-0.12, 0.25]], **f)
            self.hand_arm_default_dof_pos = torch.zeros(self.num_hand_arm_dofs, **f)
            self.hand_arm_default_dof_pos[:7] = torch.tensor([-1.571, 1.571, -0.000, 1.376, -0.000, 1.485, 2.358])
        else:   # allegro_kuka_two_arms.py:243-276
            self.target_volume_origin = torch.tensor([0, 0.0, 0.8], **f)
            self.target_volume_extent = torch.tensor([[-0.2, 0.2], [-0.5, 0.5], [-0.12, 0.25]], **f)
            dflt = torch.zeros([self.num_arms, self.num_hand_arm_dofs], **f)
            dflt[:, :7] = torch.tensor([-1.571, 1.571, -0.000, 1.6, -0.000, 1.485, 2.358])
            self.hand_arm_default_dof_pos = dflt.flatten()
        self.prev_targets = torch.zeros((N, self.num_dofs), **f)
        self.cur_targets = torch.zeros((N, self.num_dofs), **f)
        self.actions = torch.zeros((N, self.num_actions), **f)
        self.reset_goal_buf = self.reset_buf.clone()
        self.successes = torch.zeros(N, **f)
        self.prev_episode_successes = torch.zeros(N, **f)
        self.true_objective = torch.zeros(N, **f)
        self.prev_episode_true_objective = torch.zeros(N, **f)
        self.total_successes = self.total_resets = 0
        self.force_decay_t = torch.tensor(self.force_decay, **f)
        self.force_prob_range_t = torch.tensor(self.force_prob_range, **f)
        gen = torch.Generator().manual_seed(self._seed)
        u0 = torch.rand(N, generator=gen).to(dev)
        self.random_force_prob = torch.exp((torch.log(self.force_prob_range_t[0]) - torch.log(self.force_prob_range_t[1])) * u0
                                           + torch.log(self.force_prob_range_t[1]))
        if self.sim.state_layout == "soa" and self.num_arms == 1:
            # component planes like the simulator state: [n_bodies, 3, N]; rb_forces keeps the reference's [N, n_bodies, 3] indexing as a view
            self._rb_forces_planes = torch.zeros((self.num_bodies, 3, N), **f)
            self.rb_forces = self._rb_forces_planes.permute(2, 0, 1)
        else:
            self._rb_forces_planes = None
            self.rb_forces = torch.zeros((N, self.num_bodies, 3), **f)
        self.near_goal_steps = torch.zeros(N, dtype=torch.int, device=dev)
        self.lifted_object = torch.zeros(N, dtype=torch.bool, device=dev)
        self.closest_keypoint_max_dist = -torch.ones(N, **f)
        self.closest_fingertip_dist = -torch.ones([N, T] if self.num_arms == 1 else [N, self.num_arms, T], **f)
        self.furthest_hand_dist = -torch.ones([N], **f)
        self.finger_rew_coeffs = torch.ones([N, T], **f)
        # SoA: one [terms, N] block, exposed under the reference's dict keys as row views
        self._rewards_episode = torch.zeros((len(self.episode_term_keys), N), **f)
        self.rewards_episode = {k: self._rewards_episode[i] for i, k in enumerate(self.episode_term_keys)}
        self._episode_cumulative = torch.zeros((len(self.step_term_keys), N), **f)
        self.episode_cumulative = {k: self._episode_cumulative[i] for i, k in enumerate(self.step_term_keys)}
        self.last_curriculum_update = 0
        # engine-side buffers
        self.root_reset_indices = torch.zeros(2 * N, dtype=torch.int32, device=dev)
        self.hand_reset_indices = torch.zeros(N * self.num_arms, dtype=torch.int32, device=dev)
        self.reset_counts = torch.zeros(4, dtype=torch.int32, device=dev)
        self._stats = torch.zeros(C["HA_STATS_LEN"], **f)
        self._stats_pending = False
        self._tickets = torch.zeros(8, dtype=torch.int32, device=dev)
        self._object_actor_i32 = self.object_indices.to(torch.int32)
        self._goal_actor_i32 = self.goal_object_indices.to(torch.int32)
        self._hand_actor_i32 = self.arm_indices.to(torch.int32).contiguous()
        self._object_init_z = self.object_init_state[:, 2].contiguous()   # compact copy for the per-step read (task constant)
        self._dynamic = torch.zeros(8, **f)          # HaKukaDynamic (32 bytes)
        self.ext_rand: Optional[torch.Tensor] = None   # parity mode: [N, HA_KUKA_RND_STRIDE] raw draws
        self.use_fast_rng = True

    # ---- C-ABI structs -----------------------------------------------------------------------
    def _build_abi(self):
        if torch.device(self.device).type != "cuda":
            self._params = self._buffers = None      # host logic only; compute entry points need a GPU
            return
        self._lib = _lib.load()
        L = self._layout
        p = _lib.HaKukaParams()
        p.abi_version = C["HA_ABI_VERSION"]
        n_dofs = self.num_hand_arm_dofs * self.num_arms
        p.n_envs, p.n_arms, p.n_dofs, p.dofs_per_env = self.num_envs, self.num_arms, n_dofs, self.num_dofs
        p.n_bodies, p.n_tips, p.n_keypoints, p.obs_dim = (self.num_bodies, self.num_allegro_fingertips * self.num_arms, self.num_keypoints,
                                                          self.num_obs)
        p.subtask = self.subtask_id
        for a, h in enumerate(L.palm_body):
            p.palm_body[a] = h
        p.two_arm_x_ofs, p.two_arm_obj_x_ofs = 0.75, 1.1 / 2 - 0.2     # allegro_kuka_two_arms_regrasping.py:91, allegro_kuka_two_arms.py:1163-1164
        for i, t in enumerate(L.tip_body):
            p.tip_body[i] = t
        p.object_body = L.object_body
        p.max_episode_length = int(self.max_episode_length)
        p.max_consecutive_successes = int(self.max_consecutive_successes)
        p.success_steps = self.success_steps
        p.force_enabled = 1 if self.force_scale > 0.0 else 0
        for c in range(3):
            p.palm_offset[c] = float(self.palm_offset[c])
        for t in range(self.num_allegro_fingertips):
            for c in range(3):
                p.tip_offset[t][c] = float(self.fingertip_offsets[t][c])
        lo, hi = self.arm_hand_dof_lower_limits.cpu(), self.arm_hand_dof_upper_limits.cpu()
        dflt = self.hand_arm_default_dof_pos.cpu()
        for d in range(n_dofs):
            p.dof_lower[d], p.dof_upper[d], p.dof_default[d] = float(lo[d]), float(hi[d]), float(dflt[d])
            p.dof_reset_noise[d] = self.reset_dof_pos_noise_arm if d % self.num_hand_arm_dofs < 7 else self.reset_dof_pos_noise_fingers
        self._lib.ha_kuka_fill_derived(ctypes.byref(p))
        p.clamp_abs_obs = self._kernel_clamp(True)
        p.lifting_bonus_threshold, p.lifting_bonus, p.lifting_rew_scale = self.lifting_bonus_threshold, self.lifting_bonus, self.lifting_rew_scale
        p.keypoint_rew_scale = self.keypoint_rew_scale
        p.distance_delta_rew_scale = self.distance_delta_rew_scale
        p.hand_delta_rew_scale = self.distance_delta_rew_scale * 0      # "currently disabled" (:877)
        p.bonus_per_step = self.reach_goal_bonus / self.success_steps
        p.kuka_actions_penalty_scale = self.kuka_actions_penalty_scale
        p.allegro_actions_penalty_scale = self.allegro_actions_penalty_scale
        p.fall_z, p.far_tip_dist = 0.1, 1.5
        p.arm_action_gain = self.hand_dof_speed_scale * self.dt
        p.act_moving_average = self.act_moving_average
        p.one_minus_act_moving_average = 1.0 - self.act_moving_average
        p.force_decay_pow = float(torch.pow(self.force_decay_t.cpu(), self.dt / self.force_decay_interval))
        p.force_gain[0], p.force_gain[1] = float(self.object_rb_masses[0]), self.force_scale
        lr = self.force_prob_range_t.cpu()
        p.force_log_c0 = float(torch.log(lr[0]) - torch.log(lr[1]))
        p.force_log_c1 = float(torch.log(lr[1]))
        org, ext = self.target_volume_origin.cpu(), self.target_volume_extent.cpu()
        tmin, tmax = org + ext[:, 0], org + ext[:, 1]
        tsize = tmax - tmin
        for c in range(3):
            p.target_min[c], p.target_size[c] = float(tmin[c]), float(tsize[c])
        p.reset_pos_noise[0], p.reset_pos_noise[1], p.reset_pos_noise[2] = (self.reset_position_noise_x, self.reset_position_noise_y,
                                                                          self.reset_position_noise_z)
        p.reset_dof_vel_noise = self.reset_dof_vel_noise
        # regular Isaac Gym actor numbering (row = env * actors_per_env + slot) lets the gather skip the index tables
        ar = torch.arange(self.num_envs, dtype=torch.long, device=self.device) * L.actors_per_env
        regular = bool(torch.equal(self.object_indices, ar + L.object_actor)) and bool(torch.equal(self.goal_object_indices, ar + L.goal_actor))
        p.root_regular, p.actors_per_env, p.object_slot, p.goal_slot = int(regular), L.actors_per_env, L.object_actor, L.goal_actor
        p.state_layout = C["HA_STATE_SOA"] if self.sim.state_layout == "soa" else C["HA_STATE_ISAACGYM"]
        self._params = p
        self._observe_params = _lib.HaKukaParams.from_buffer_copy(p)
        self._observe_params.clamp_abs_obs = 0.0    # VecTask.reset() clips on the host side copy only
        self._stats_deferred = self.num_arms == 1 or (self.num_envs + 31) // 32 <= C["HA_KUKA_STAT_SLOTS"]
        if self._stats_deferred:   # one slot of partials per CTA of the post kernel, folded by the next pre step / finalize_stats()
            init = torch.tensor([int(self._lib.ha_kuka_stat_accum_init(i)) for i in range(C["HA_KUKA_N_PARTIAL_STATS"])], dtype=torch.int64)
            self._stat_accum = init[:, None].repeat(1, C["HA_KUKA_STAT_SLOTS"]).contiguous().to(self.device)
        else:
            self._stat_accum = torch.tensor([int(self._lib.ha_kuka_stat_accum_init(i)) for i in range(C["HA_STATS_LEN"])],
                                            dtype=torch.int64, device=self.device)
        self._buffers = _lib.HaKukaBuffers()
        self._bind_buffers()
        self._write_dynamic()

    def _bind_buffers(self):
        b = self._buffers
        ptr = lambda t: ctypes.c_void_p(t.data_ptr())
        self._bound = dict(
            root_state=self.sim.root_state, rigid_body_state=self.sim.rigid_body_state, dof_state=self.sim.dof_state,
            object_actor=self._object_actor_i32, goal_actor=self._goal_actor_i32, hand_actor=self._hand_actor_i32,
            object_init_state=self.object_init_state, object_init_z=self._object_init_z, object_keypoint_offsets=self.object_keypoint_offsets,
            object_scales=self.object_scales, finger_rew_coeffs=self.finger_rew_coeffs, goal_states=self.goal_states,
            obs_buf=self.obs_buf, rew_buf=self.rew_buf, reset_buf=self.reset_buf, reset_goal_buf=self.reset_goal_buf,
            progress_buf=self.progress_buf, randomize_buf=self.randomize_buf, timeout_buf=self.timeout_buf, successes=self.successes,
            prev_episode_successes=self.prev_episode_successes, true_objective=self.true_objective,
            prev_episode_true_objective=self.prev_episode_true_objective, near_goal_steps=self.near_goal_steps,
            lifted_object=self.lifted_object, closest_keypoint_max_dist=self.closest_keypoint_max_dist,
            closest_fingertip_dist=self.closest_fingertip_dist, furthest_hand_dist=self.furthest_hand_dist,
            rewards_episode=self._rewards_episode, episode_cumulative=self._episode_cumulative, prev_targets=self.prev_targets,
            cur_targets=self.cur_targets, actions=self.actions,
            rb_forces=self.rb_forces if self._rb_forces_planes is None else self._rb_forces_planes, random_force_prob=self.random_force_prob,
            root_reset_indices=self.root_reset_indices, hand_reset_indices=self.hand_reset_indices, reset_counts=self.reset_counts,
            stats=self._stats, stat_accum=self._stat_accum, tickets=self._tickets, dynamic=self._dynamic)
        for name, t in self._bound.items():
            assert t.is_contiguous(), name
            setattr(b, name, t.data_ptr())
        self._bound_cursor = self.sim.cursor

    def _rebind_sim_state(self):
        """Point the engine at the simulator's current tensors (snapshot ring = pointer swap)."""
        if self._bound_cursor != self.sim.cursor:
            b = self._buffers
            b.root_state, b.rigid_body_state, b.dof_state = (self.sim.root_state.data_ptr(), self.sim.rigid_body_state.data_ptr(),
                                                             self.sim.dof_state.data_ptr())
            self._bound_cursor = self.sim.cursor

    def _write_dynamic(self, seed: Optional[int] = None):
        scale, offset = true_objective_coeffs(self.success_tolerance, self.initial_tolerance, self.target_tolerance)
        host = torch.tensor([self.success_tolerance * self.keypoint_scale, scale, offset], dtype=torch.float32)
        self._dynamic[0:3].copy_(host, non_blocking=True)
        if seed is not None:
            self._dynamic.view(torch.int64)[2:4] = torch.tensor([seed, 0], dtype=torch.int64)
        self._dynamic_tol = self.success_tolerance

    def seed_rng(self, seed: int):
        """Seed of the in-kernel Philox stream (fast mode); the step counter restarts at 0."""
        self._write_dynamic(seed=int(seed))

    def _kernel_clamp(self, with_vec_task_clip: bool) -> float:
        clamp = self.clamp_abs_observations if self.clamp_abs_observations > 0 else math.inf
        if with_vec_task_clip:
            clamp = min(clamp, self.clip_obs)        # clamp_obs (:1174-1176) then the VecTask clip (vec_task.py:432)
        return clamp if math.isfinite(clamp) else 0.0

    def _set_in_kernel_obs_clip(self, on: bool) -> None:
        if self._params is not None:
            self._params.clamp_abs_obs = self._kernel_clamp(on)

    def set_post_grid(self, ctas: int):
        """Test / tuning hook (HaKukaParams.post_grid): force the CTA count of the persistent post grid; 0 restores one full wave.
        A small grid makes a small env block walk many tiles per CTA, the regime of the 32768-env headline size."""
        self._params.post_grid = int(ctas)
        self._observe_params.post_grid = int(ctas)

    def _stream(self) -> ctypes.c_void_p:
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _require_engine(self):
        if self._params is None:
            raise _lib.HandArmLibraryError("the task hot path runs only on a CUDA device (no CPU fallback); "
                                           f"this task was created on device {self.device!r}")

    # ---- properties mirroring reference views --------------------------------------------------
    @property
    def root_state_tensor(self):
        return self.sim.root_state if self.sim.state_layout == "isaacgym" else self.sim.root_rows()

    @property
    def rigid_body_states(self):
        return self.sim.body_rows()

    @property
    def dof_state(self):
        return self.sim.dof_state

    @property
    def arm_hand_dof_pos(self):
        return self.sim.dof_state.view(self.num_envs, -1, 2)[:, : self.num_hand_arm_dofs * self.num_arms, 0]

    @property
    def arm_hand_dof_vel(self):
        return self.sim.dof_state.view(self.num_envs, -1, 2)[:, : self.num_hand_arm_dofs * self.num_arms, 1]

    # ---- env state side channel (allegro_kuka_base.py:472-493) ------------------------------------
    def get_env_state(self):
        return dict(success_tolerance=self.success_tolerance)

    def set_env_state(self, env_state):
        if env_state is None:
            return
        value = env_state.get("success_tolerance", None)
        if value is not None:
            self.success_tolerance = value
            if self._params is not None:
                self._write_dynamic()

    # ---- the hot path --------------------------------------------------------------------------
    def _rand_ptr(self):
        if self.ext_rand is not None:
            stride = (25 if self.num_arms == 1 else 34) + 2 * self.num_hand_arm_dofs * self.num_arms   # HA_KUKA_RND_STRIDE / HA_KUKA2_RND_STRIDE
            assert self.ext_rand.shape == (self.num_envs, stride) and self.ext_rand.dtype == torch.float32 and self.ext_rand.is_contiguous()
            self._params.use_fast_rng = 0
            return ctypes.c_void_p(self.ext_rand.data_ptr())
        self._params.use_fast_rng = 1
        return ctypes.c_void_p(0)

    # ---- loadInitialStates (allegro_kuka_base.py:1545-1592 file, :1292-1313 use; off in the shipped configs) ----
    def load_initial_states(self):
        from .. import state_io
        root, dof = state_io.load_initial_states(self.load_states_filename)
        self.initial_root_state_tensors = root.to(self.device, torch.float32)
        self.initial_dof_state_tensors = dof.to(self.device, torch.float32)
        self.num_initial_states = int(root.shape[0])
        print(f"{self.num_initial_states} states loaded from file {self.load_states_filename}!")

    def _initial_state_ids(self) -> Optional[torch.Tensor]:
        """The reset envs in ascending order, read BEFORE the kernel consumes the mask.  The recorded states are handed out
        consecutively in that order, which needs the ids on the host exactly like the reference's ``nonzero()``: the one place
        where this option costs a sync (and why it cannot be captured in a CUDA graph)."""
        if not self.should_load_initial_states:
            return None
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError("loadInitialStates reads the reset ids on the host (allegro_kuka_base.py:1292-1313): not capturable")
        return self.reset_buf.nonzero(as_tuple=False).squeeze(-1)

    def _apply_initial_states(self, env_ids: torch.Tensor) -> None:
        """allegro_kuka_base.py:1292-1313, after the kernel's re-randomisation: every DOF state and the object's root row of the reset
        envs come from the file; prev / cur targets keep the re-randomised position."""
        k = int(env_ids.numel())
        if k == 0:
            return
        if k > self.num_initial_states:
            print(f"Not enough initial states to load {k}/{self.num_initial_states}...")
            return
        if self.initial_state_idx + k > self.num_initial_states:
            self.initial_state_idx = 0
        i0 = self.initial_state_idx
        self.sim.dof_state.view(self.num_envs, -1, 2)[env_ids] = self.initial_dof_state_tensors[i0:i0 + k]
        slot = self.sim.layout.object_actor                # cube_object_idx = self.object_indices[0] (:1308)
        self.sim.root_rows()[env_ids, slot] = self.initial_root_state_tensors[i0:i0 + k, slot]
        self.initial_state_idx += k

    def pre_physics_step(self, actions: torch.Tensor):
        """allegro_kuka_base.py:1355-1424 as one kernel (masks consumed in place, no nonzero()/unique())."""
        self._require_engine()
        if actions.dtype != torch.float32 or not actions.is_contiguous():
            actions = actions.float().contiguous()
        self._rebind_sim_state()
        rnd = self._rand_ptr()
        load_ids = self._initial_state_ids()
        _lib.check(self._lib.ha_kuka_pre_step(ctypes.byref(self._params), ctypes.byref(self._buffers), ctypes.c_void_p(actions.data_ptr()),
                                              rnd, self._stream()))
        if load_ids is not None:
            self._apply_initial_states(load_ids)
        self._stats_pending = False       # the pre launch folded the previous step's statistic slots
        # what the reference hands to PhysX after this point (recorded, not executed):
        sim = self.sim
        if sim.record_calls:
            sim.record("set_dof_position_target_tensor_indexed", self.prev_targets, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_dof_state_tensor_indexed", sim.dof_state, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_actor_root_state_tensor_indexed", sim.root_state, self.root_reset_indices, self.reset_counts[0:1])
            sim.record("set_dof_position_target_tensor", self.cur_targets)
            if self.force_scale > 0.0:
                sim.record("apply_rigid_body_force_tensors", self.rb_forces, None, "LOCAL_SPACE")

    def _extra_curriculum(self):
        pass

    def advance_frames(self, n: int) -> None:
        """Host-side frame accounting for steps that were replayed from a CUDA graph (the kernels ran,
        the Python hooks did not): keeps ``frame_since_restart`` and the tolerance curriculum going."""
        self.frame_since_restart += n
        self.control_steps += n
        self._extra_curriculum()

    def post_physics_step(self):
        """allegro_kuka_base.py:1426-1443 + vec_task.py:424-432 as one kernel."""
        self._require_engine()
        self.frame_since_restart += 1
        self._extra_curriculum()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_kuka_post_step(ctypes.byref(self._params), ctypes.byref(self._buffers), self._stream()))
        self._stats_pending = self._stats_deferred
        # 0-dim views of the statistics vector: filled in place by the next pre step or by finalize_stats() / the `stats` property
        ex, st = self.extras, self._stats
        ex["successes"] = st[C["HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES"]]
        ex["true_objective"] = self.true_objective
        ex["true_objective_mean"] = st[C["HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE"]]
        ex["true_objective_min"] = st[C["HA_KUKA_STAT_MIN_TRUE_OBJECTIVE"]]
        ex["true_objective_max"] = st[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]]
        ex["rewards_episode"] = self.rewards_episode
        ex["episode_cumulative"] = self.episode_cumulative
        ex["scalars"] = {"success_tolerance": self.success_tolerance}

    def finalize_stats(self) -> None:
        """Fold the post kernel's per-CTA statistic slots into the statistics vector now (one small launch) instead of at
        the head of the next ``pre_physics_step``.  Nothing pending (or two arms beyond the slot count): no-op."""
        if self._stats_pending:
            _lib.check(self._lib.ha_kuka_stats_finalize(ctypes.byref(self._params), ctypes.byref(self._buffers), self._stream()))
            self._stats_pending = False

    @property
    def stats(self) -> torch.Tensor:
        """Episode statistics of the last post step ([HA_STATS_LEN] float32, ``HA_KUKA_STAT_*``)."""
        self.finalize_stats()
        return self._stats

    @property
    def stats_folded(self) -> torch.Tensor:
        """The statistics vector as the last pre step (or ``finalize_stats``) left it: no launch, one post step old inside a
        step loop.  What a per-epoch consumer (logging, cross-rank exchange) reads without touching the stream."""
        return self._stats

    def compute_observations(self):
        """allegro_kuka_base.py:991-1089 only (used by VecTask.reset()): obs + lazy init, no reward."""
        self._require_engine()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_kuka_observe(ctypes.byref(self._observe_params), ctypes.byref(self._buffers), self._stream()))
        return self.obs_buf, self.full_state_size - 1

    def reset_idx(self, env_ids: torch.Tensor) -> None:
        """allegro_kuka_base.py:1246-1353 for an explicit id list (reset_done path)."""
        self._require_engine()
        saved = self.reset_buf.clone()
        self.reset_buf.zero_()
        self.reset_buf[env_ids] = 1
        self._rebind_sim_state()
        _lib.check(self._lib.ha_kuka_reset_idx(ctypes.byref(self._params), ctypes.byref(self._buffers), self._rand_ptr(), self._stream()))
        if self.should_load_initial_states:
            self._apply_initial_states(torch.as_tensor(env_ids, device=self.device).reshape(-1))      # in the caller's order, as the reference
        self._stats_pending = False
        saved[env_ids] = 0
        self.reset_buf.copy_(saved)
        self.extras["scalars"] = {"success_tolerance": self.success_tolerance}

    def get_state(self):
        # states_buf is never written by this task family (it stays zero in the reference as well)
        return self.states_buf if self.rl_device == self.device else self.states_buf.to(self.rl_device)

    def bytes_per_env_step(self) -> int:
        return int(self._lib.ha_kuka_bytes_per_env_step(ctypes.byref(self._params)))


class AllegroKukaReorientation(AllegroKukaBase):
    """allegro_kuka_reorientation.py:41-149."""
    subtask_id = C["HA_KUKA_REORIENTATION"]

    def _object_keypoint_offsets(self):
        return [[1, 1, 1], [1, 1, -1], [-1, -1, 1], [-1, -1, -1]]

    def _extra_curriculum(self):
        if torch.device(self.device).type == "cuda" and torch.cuda.is_current_stream_capturing():
            return      # the rule reads a device mean on the host: it waits for the next eager step / advance_frames()
        tol, self.last_curriculum_update = tolerance_curriculum(
            self.last_curriculum_update, self.frame_since_restart, self.tolerance_curriculum_interval, self.prev_episode_successes,
            self.success_tolerance, self.initial_tolerance, self.target_tolerance, self.tolerance_curriculum_increment)
        if tol != self.success_tolerance:
            self.success_tolerance = tol
            self._write_dynamic()


class AllegroKukaRegrasping(AllegroKukaReorientation):
    """allegro_kuka_regrasping.py:40-123: grasp, lift and hold the object at a target position.  One keypoint (the object centre,
    99-wide observation); a new goal (`_reset_target` :76-100) draws a position only and puts the object back on the table
    (reset_object_pose, lifted_object = False); no far-fingertip reset rule; same tolerance curriculum and true objective."""
    subtask_id = C["HA_KUKA_REGRASPING"]

    def _object_keypoint_offsets(self):
        return [[0, 0, 0]]


class AllegroKukaThrow(AllegroKukaRegrasping):
    """allegro_kuka_throw.py:40-124: throw the object into a bucket.  The bucket takes the goal actor's slot; `_reset_target`
    (:90-117) places it left / right of the table (4 draws), the goal is 5 cm above it, the object goes back to the table.  The
    class does not override `_extra_curriculum` (allegro_kuka_base.py:468-469: no curriculum)."""
    subtask_id = C["HA_KUKA_THROW"]

    def _extra_curriculum(self):
        pass


class AllegroKukaTwoArmsRegrasping(AllegroKukaReorientation):
    """allegro_kuka_two_arms.py (``AllegroKukaTwoArmsBase``) + allegro_kuka_two_arms_regrasping.py:40-135: two Kuka-Allegro arms pass
    the object between them; 46 DOFs all position-controlled, one keypoint (the object centre), 169-wide observation."""
    subtask_id = C["HA_KUKA_REGRASPING"]
    num_arms = 2
    default_layout = KUKA2_LAYOUT
    episode_term_keys, step_term_keys = EPISODE_TERM_KEYS_2ARMS, STEP_TERM_KEYS_2ARMS

    def _object_keypoint_offsets(self):
        return [[0, 0, 0]]


class AllegroKukaTwoArmsReorientation(AllegroKukaTwoArmsRegrasping):
    """allegro_kuka_two_arms_reorientation.py:18-158: the two-arm task with a goal POSE -- four keypoints (:25-31), 196-wide
    observation, ``_reset_target`` draws a position, a side and a rotation and leaves the object where it is (:97-137); tolerance
    curriculum and true objective as in the single-arm reorientation (:142-158)."""
    subtask_id = C["HA_KUKA_REORIENTATION"]

    def _object_keypoint_offsets(self):
        return [[1, 1, 1], [1, 1, -1], [-1, -1, 1], [-1, -1, -1]]
