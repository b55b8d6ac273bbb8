"""ShadowHand and AllegroHand in-hand reorientation on the fused sm_100a kernels.

Drop-in for ``isaacgymenvs/tasks/shadow_hand.py`` (``ShadowHand``) and ``allegro_hand.py`` (``AllegroHand``): same
constructor, config keys, attribute names (the reference's ``shadow_hand_*`` names are kept for both hands, as the
reference itself does), buffers and ``extras`` keys.  ``pre_physics_step`` / ``post_physics_step`` are one kernel launch
each (``ha_hand_pre_step`` / ``ha_hand_post_step``); nothing on the step path synchronises with the host.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional

import numpy as np
import torch

from .. import _lib
from ..sim import (ALLEGRO_HAND_LAYOUT, ALLEGRO_HAND_LOWER, ALLEGRO_HAND_UPPER, SHADOW_HAND_ACTUATED, SHADOW_HAND_LAYOUT,
                   SHADOW_HAND_LOWER, SHADOW_HAND_UPPER, ActorLayout, SyntheticSim)
from ..vec_task import VecTask

C = _lib.CONSTS
OBS_TYPES = {"openai": C["HA_HAND_OBS_OPENAI"], "full_no_vel": C["HA_HAND_OBS_FULL_NO_VEL"], "full": C["HA_HAND_OBS_FULL"],
             "full_state": C["HA_HAND_OBS_FULL_STATE"]}


class _HandTask(VecTask):
    """Shared body of ``ShadowHand`` (shadow_hand.py:42-200) and ``AllegroHand`` (allegro_hand.py:41-190)."""
    is_shadow = True
    num_obs_dict = {}
    default_layout: ActorLayout = SHADOW_HAND_LAYOUT
    dof_lower, dof_upper = SHADOW_HAND_LOWER, SHADOW_HAND_UPPER
    actuated = SHADOW_HAND_ACTUATED
    object_start = (0.0, -0.39, 0.54)      # hand pose (0, 0, 0.5) + (0, -0.39, 0.04)  (shadow_hand.py:305-316)

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture=False, force_render=False,
                 sim: Optional[SyntheticSim] = None, layout: Optional[ActorLayout] = None, seed: int = 1234, snapshots: int = 1):
        self.cfg = cfg
        env = cfg["env"]
        self._injected_sim, self._layout, self._seed, self._snapshots = sim, layout or self.default_layout, seed, snapshots
        self.randomize = cfg["task"]["randomize"]
        self.randomization_params = cfg["task"]["randomization_params"]
        self.aggregate_mode = env["aggregateMode"]
        self.dist_reward_scale = env["distRewardScale"]
        self.rot_reward_scale = env["rotRewardScale"]
        self.action_penalty_scale = env["actionPenaltyScale"]
        self.success_tolerance = env["successTolerance"]
        self.reach_goal_bonus = env["reachGoalBonus"]
        self.fall_dist = env["fallDistance"]
        self.fall_penalty = env["fallPenalty"]
        self.rot_eps = env["rotEps"]
        self.vel_obs_scale = 0.2
        self.force_torque_obs_scale = 10.0
        self.reset_position_noise = env["resetPositionNoise"]
        self.reset_rotation_noise = env["resetRotationNoise"]
        self.reset_dof_pos_noise = env["resetDofPosRandomInterval"]
        self.reset_dof_vel_noise = env["resetDofVelRandomInterval"]
        self.force_scale = env.get("forceScale", 0.0)
        self.force_prob_range = env.get("forceProbRange", [0.001, 0.1])
        self.force_decay = env.get("forceDecay", 0.99)
        self.force_decay_interval = env.get("forceDecayInterval", 0.08)
        self.shadow_hand_dof_speed_scale = env["dofSpeedScale"]
        self.use_relative_control = env["useRelativeControl"]
        self.act_moving_average = env["actionsMovingAverage"]
        self.debug_viz = False
        self.max_episode_length = env["episodeLength"]
        self.reset_time = env.get("resetTime", -1.0)
        self.print_success_stat = env["printNumSuccesses"]
        self.max_consecutive_successes = env["maxConsecutiveSuccesses"]
        self.av_factor = env.get("averFactor", 0.1)
        self.object_type = env["objectType"]
        assert self.object_type in ["block", "egg", "pen"]
        self.ignore_z = self.object_type == "pen"
        self.obs_type = env["observationType"]
        if self.obs_type not in self.num_obs_dict:
            raise Exception("Unknown type of observations!\nobservationType should be one of: " + str(list(self.num_obs_dict)))
        self.up_axis = "z"
        self.num_fingertips = len(self._layout.tip_body) if self.is_shadow else 0
        self.asymmetric_obs = env["asymmetric_observations"]
        num_states = self.num_obs_dict["full_state"] if self.asymmetric_obs else 0
        cfg["env"]["numObservations"] = self.num_obs_dict[self.obs_type]
        cfg["env"]["numStates"] = num_states
        cfg["env"]["numActions"] = len(self.actuated)
        super().__init__(cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture, force_render)
        self.dt = self.sim_params_dt
        if self.reset_time > 0.0:
            self.max_episode_length = int(round(self.reset_time / (env.get("controlFrequencyInv", 1) * self.dt)))
        self._allocate_task_state()
        self._build_abi()

    # ---- construction ------------------------------------------------------------------------
    def create_sim(self):
        """Replaces asset loading / actor creation (shadow_hand.py:202-413) by the synthetic producer."""
        L = self._layout
        self.up_axis_idx = 2
        self.sim = self._injected_sim or SyntheticSim(self.num_envs, L, self.device, seed=self._seed, snapshots=self._snapshots)
        self.gym = self.sim
        N, dev = self.num_envs, self.device
        self.num_shadow_hand_dofs = len(self.dof_lower)
        self.num_bodies, self.num_dofs = L.bodies_per_env, L.dofs_per_env
        self.shadow_hand_dof_lower_limits = torch.tensor(self.dof_lower, dtype=torch.float, device=dev)
        self.shadow_hand_dof_upper_limits = torch.tensor(self.dof_upper, dtype=torch.float, device=dev)
        self.shadow_hand_dof_default_pos = torch.zeros(self.num_shadow_hand_dofs, dtype=torch.float, device=dev)
        self.shadow_hand_dof_default_vel = torch.zeros(self.num_shadow_hand_dofs, dtype=torch.float, device=dev)
        self.actuated_dof_indices = torch.tensor(self.actuated, dtype=torch.long, device=dev)
        self.fingertip_handles = torch.tensor(list(L.tip_body), dtype=torch.long, device=dev)
        self.object_rb_handles = torch.tensor([L.object_body], dtype=torch.long, device=dev)
        self.object_rb_masses = torch.tensor([0.1], dtype=torch.float, device=dev)
        ar = torch.arange(N, dtype=torch.long, device=dev) * L.actors_per_env
        self.hand_indices, self.object_indices, self.goal_object_indices = ar + L.hand_actor[0], ar + L.object_actor, ar + L.goal_actor
        init = torch.zeros(N, 13, dtype=torch.float)
        init[:, 0:3] = torch.tensor(self.object_start)
        init[:, 6] = 1.0
        self.object_init_state = init.to(dev)
        self.goal_displacement_tensor = torch.tensor([-0.2, -0.06, 0.12], dtype=torch.float, device=dev)
        self.goal_states = self.object_init_state.clone()
        self.goal_states[:, self.up_axis_idx] -= 0.04
        self.goal_init_state = self.goal_states.clone()

    def _allocate_task_state(self):
        """shadow_hand.py:160-200, same names and dtypes."""
        N, dev = self.num_envs, self.device
        f = dict(dtype=torch.float, device=dev)
        self.shadow_hand_default_dof_pos = torch.zeros(self.num_shadow_hand_dofs, **f)
        self.prev_targets = torch.zeros((N, self.num_dofs), **f)
        self.cur_targets = torch.zeros((N, self.num_dofs), **f)
        self.actions = torch.zeros((N, self.num_actions), **f)
        self.x_unit_tensor = torch.tensor([1, 0, 0], **f).repeat((N, 1))
        self.y_unit_tensor = torch.tensor([0, 1, 0], **f).repeat((N, 1))
        self.z_unit_tensor = torch.tensor([0, 0, 1], **f).repeat((N, 1))
        self.reset_goal_buf = self.reset_buf.clone()
        self.successes = torch.zeros(N, **f)
        self._consecutive_successes = torch.zeros(1, **f)
        self.total_successes = self.total_resets = 0
        self.force_decay_t = torch.tensor(self.force_decay, **f)
        self.force_prob_range_t = torch.tensor(self.force_prob_range, **f)
        gen = torch.Generator().manual_seed(self._seed)
        u0 = torch.rand(N, generator=gen).to(dev)
        self.random_force_prob = torch.exp((torch.log(self.force_prob_range_t[0]) - torch.log(self.force_prob_range_t[1])) * u0
                                           + torch.log(self.force_prob_range_t[1]))
        self.rb_forces = torch.zeros((N, self.num_bodies, 3), **f)
        # engine-side buffers
        self.root_reset_indices = torch.zeros(2 * N, dtype=torch.int32, device=dev)
        self.hand_reset_indices = torch.zeros(N, dtype=torch.int32, device=dev)
        self.reset_counts = torch.zeros(4, dtype=torch.int32, device=dev)
        self._stats = torch.zeros(C["HA_STATS_LEN"], **f)
        # one slot of integer partials per CTA of the post kernel, folded by the next pre step / finalize_stats()
        self._stat_accum = torch.zeros(C["HA_HAND_N_PARTIAL_STATS"] * C["HA_HAND_STAT_SLOTS"], dtype=torch.int64, device=dev)
        n_tiles = (self.num_envs + 31) // 32
        self._stats_deferred = n_tiles <= C["HA_HAND_STAT_SLOTS"]
        self._stats_pending = False
        self._tickets = torch.zeros(8, dtype=torch.int32, device=dev)
        self._object_actor_i32 = self.object_indices.to(torch.int32)
        self._goal_actor_i32 = self.goal_object_indices.to(torch.int32)
        self._hand_actor_i32 = self.hand_indices.to(torch.int32)
        self._dynamic = torch.zeros(2, dtype=torch.int64, device=dev)      # HaHandDynamic: rng_seed, rng_step
        self.ext_rand: Optional[torch.Tensor] = None                       # parity mode: [N, HA_HAND_RND_STRIDE] raw draws

    # ---- C-ABI structs -----------------------------------------------------------------------
    def _build_abi(self):
        if torch.device(self.device).type != "cuda":
            self._params = self._buffers = None      # host logic only; compute entry points need a GPU
            return
        if self.sim.state_layout != "isaacgym":
            raise NotImplementedError("the in-hand kernels read Isaac Gym's AoS state tensors (HA_STATE_ISAACGYM)")
        self._lib = _lib.load()
        L = self._layout
        p = _lib.HaHandParams()
        p.abi_version = C["HA_ABI_VERSION"]
        p.n_envs, p.is_shadow = self.num_envs, int(self.is_shadow)
        p.n_dofs, p.dofs_per_env, p.n_actions = self.num_shadow_hand_dofs, self.num_dofs, self.num_actions
        p.n_bodies, p.n_tips = self.num_bodies, self.num_fingertips
        p.obs_type, p.obs_dim, p.states_dim = OBS_TYPES[self.obs_type], self.num_obs, self.num_states
        p.object_body = L.object_body
        p.max_episode_length = int(self.max_episode_length)
        p.max_consecutive_successes = int(self.max_consecutive_successes)
        p.use_relative_control = int(bool(self.use_relative_control))
        p.ignore_z_rot = int(self.ignore_z)
        p.force_enabled = 1 if self.force_scale > 0.0 else 0
        ar = torch.arange(self.num_envs, dtype=torch.long, device=self.device) * L.actors_per_env
        regular = bool(torch.equal(self.object_indices, ar + L.object_actor)) and bool(torch.equal(self.goal_object_indices, ar + L.goal_actor))
        p.root_regular, p.actors_per_env, p.object_slot, p.goal_slot = int(regular), L.actors_per_env, L.object_actor, L.goal_actor
        p.up_axis_idx = self.up_axis_idx
        for a, d in enumerate(self.actuated):
            p.actuated_dof[a] = d
        for t, bd in enumerate(L.tip_body[: self.num_fingertips]):
            p.tip_body[t] = bd
        lo, hi = self.shadow_hand_dof_lower_limits.cpu(), self.shadow_hand_dof_upper_limits.cpu()
        for d in range(self.num_shadow_hand_dofs):
            p.dof_lower[d], p.dof_upper[d], p.dof_default[d] = float(lo[d]), float(hi[d]), 0.0
        self._lib.ha_hand_fill_derived(ctypes.byref(p))
        p.clip_obs = self.clip_obs if math.isfinite(self.clip_obs) else 0.0
        p.vel_obs_scale, p.force_torque_obs_scale = self.vel_obs_scale, self.force_torque_obs_scale
        p.dist_reward_scale, p.rot_reward_scale, p.rot_eps = self.dist_reward_scale, self.rot_reward_scale, self.rot_eps
        p.action_penalty_scale, p.success_tolerance = self.action_penalty_scale, self.success_tolerance
        p.reach_goal_bonus, p.fall_dist, p.fall_penalty, p.av_factor = self.reach_goal_bonus, self.fall_dist, self.fall_penalty, self.av_factor
        p.rel_action_gain = self.shadow_hand_dof_speed_scale * self.dt
        p.act_moving_average = self.act_moving_average
        p.one_minus_act_moving_average = 1.0 - self.act_moving_average
        p.force_decay_pow = float(torch.pow(self.force_decay_t.cpu(), self.dt / self.force_decay_interval))
        p.force_gain[0], p.force_gain[1] = float(self.object_rb_masses[0]), self.force_scale
        lr = self.force_prob_range_t.cpu()
        p.force_log_c0 = float(torch.log(lr[0]) - torch.log(lr[1]))
        p.force_log_c1 = float(torch.log(lr[1]))
        p.reset_position_noise, p.reset_dof_pos_noise, p.reset_dof_vel_noise = (self.reset_position_noise, self.reset_dof_pos_noise,
                                                                                 self.reset_dof_vel_noise)
        gd = self.goal_displacement_tensor.cpu()
        for c in range(3):
            p.goal_displacement[c] = float(gd[c])
        self._params = p
        self._observe_params = _lib.HaHandParams.from_buffer_copy(p)
        self._observe_params.clip_obs = 0.0          # VecTask.reset() clips the copy it returns, not obs_buf
        self._buffers = _lib.HaHandBuffers()
        self._bind_buffers()

    def _bind_buffers(self):
        b = self._buffers
        self._bound = dict(
            root_state=self.sim.root_state, rigid_body_state=self.sim.rigid_body_state, dof_state=self.sim.dof_state,
            dof_force=self.sim.dof_force, force_sensors=self.sim.force_sensors, object_actor=self._object_actor_i32,
            goal_actor=self._goal_actor_i32, hand_actor=self._hand_actor_i32, object_init_state=self.object_init_state,
            goal_init_state=self.goal_init_state, goal_states=self.goal_states, obs_buf=self.obs_buf, rew_buf=self.rew_buf,
            reset_buf=self.reset_buf, reset_goal_buf=self.reset_goal_buf, progress_buf=self.progress_buf, randomize_buf=self.randomize_buf,
            timeout_buf=self.timeout_buf, successes=self.successes, consecutive_successes=self._consecutive_successes,
            prev_targets=self.prev_targets, cur_targets=self.cur_targets, actions=self.actions, rb_forces=self.rb_forces,
            random_force_prob=self.random_force_prob, root_reset_indices=self.root_reset_indices,
            hand_reset_indices=self.hand_reset_indices, reset_counts=self.reset_counts, stats=self._stats, stat_accum=self._stat_accum,
            tickets=self._tickets, dynamic=self._dynamic)
        if self.num_states > 0:
            self._bound["states_buf"] = self.states_buf
        for name, t in self._bound.items():
            assert t.is_contiguous(), name
            setattr(b, name, t.data_ptr())
        self._bound_cursor = self.sim.cursor

    def _rebind_sim_state(self):
        if self._bound_cursor != self.sim.cursor:
            b, s = self._buffers, self.sim
            b.root_state, b.rigid_body_state, b.dof_state = s.root_state.data_ptr(), s.rigid_body_state.data_ptr(), s.dof_state.data_ptr()
            b.dof_force, b.force_sensors = s.dof_force.data_ptr(), s.force_sensors.data_ptr()
            self._bound_cursor = s.cursor

    def seed_rng(self, seed: int):
        """Seed of the in-kernel Philox stream (fast mode); the step counter restarts at 0."""
        self._dynamic.copy_(torch.tensor([int(seed), 0], dtype=torch.int64))

    def _set_in_kernel_obs_clip(self, on: bool) -> None:
        if self._params is not None:
            self._params.clip_obs = self.clip_obs if (on and math.isfinite(self.clip_obs)) else 0.0

    def _stream(self) -> ctypes.c_void_p:
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _require_engine(self):
        if self._params is None:
            raise _lib.HandArmLibraryError("the task hot path runs only on a CUDA device (no CPU fallback); "
                                           f"this task was created on device {self.device!r}")

    # ---- reference views ---------------------------------------------------------------------------
    @property
    def root_state_tensor(self):
        return self.sim.root_state

    @property
    def rigid_body_states(self):
        return self.sim.body_rows()

    @property
    def dof_state(self):
        return self.sim.dof_state

    @property
    def shadow_hand_dof_pos(self):
        return self.sim.dof_state.view(self.num_envs, -1, 2)[:, : self.num_shadow_hand_dofs, 0]

    @property
    def shadow_hand_dof_vel(self):
        return self.sim.dof_state.view(self.num_envs, -1, 2)[:, : self.num_shadow_hand_dofs, 1]

    @property
    def dof_force_tensor(self):
        return self.sim.dof_force

    @property
    def vec_sensor_tensor(self):
        return self.sim.force_sensors

    # ---- the hot path --------------------------------------------------------------------------
    def _rand_ptr(self):
        if self.ext_rand is not None:
            stride = 18 + 2 * self.num_shadow_hand_dofs        # HA_HAND_RND_STRIDE
            assert self.ext_rand.shape == (self.num_envs, stride) and self.ext_rand.dtype == torch.float32 and self.ext_rand.is_contiguous()
            self._params.use_fast_rng = 0
            return ctypes.c_void_p(self.ext_rand.data_ptr())
        self._params.use_fast_rng = 1
        return ctypes.c_void_p(0)

    def pre_physics_step(self, actions: torch.Tensor):
        """shadow_hand.py:670-708 / allegro_hand.py:586-625 as one kernel."""
        self._require_engine()
        if actions.dtype != torch.float32 or not actions.is_contiguous():
            actions = actions.float().contiguous()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_hand_pre_step(ctypes.byref(self._params), ctypes.byref(self._buffers), ctypes.c_void_p(actions.data_ptr()),
                                              self._rand_ptr(), self._stream()))
        self._stats_pending = False       # the pre launch folded the previous step's statistic slots
        sim = self.sim
        if sim.record_calls:
            sim.record("set_actor_root_state_tensor_indexed", sim.root_state, self.root_reset_indices, self.reset_counts[0:1])
            sim.record("set_dof_position_target_tensor_indexed", self.prev_targets, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_dof_state_tensor_indexed", sim.dof_state, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_dof_position_target_tensor", self.cur_targets)
            if self.force_scale > 0.0:
                sim.record("apply_rigid_body_force_tensors", self.rb_forces, None, "LOCAL_SPACE")

    def post_physics_step(self):
        """post_physics_step + compute_reward wrapper (shadow_hand.py:415-435) + vec_task.py:424-432 as one kernel."""
        self._require_engine()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_hand_post_step(ctypes.byref(self._params), ctypes.byref(self._buffers), self._stream()))
        self._stats_pending = self._stats_deferred
        # mean of the [1] tensor: a 0-dim view, no launch; filled in place by the next pre step or by finalize_stats()
        self.extras["consecutive_successes"] = self._consecutive_successes[0]

    def finalize_stats(self) -> None:
        """Fold the post kernel's per-CTA statistic slots into ``stats`` / ``consecutive_successes`` now (one small launch)
        instead of at the head of the next ``pre_physics_step``.  Nothing pending: no-op."""
        if self._stats_pending:
            _lib.check(self._lib.ha_hand_stats_finalize(ctypes.byref(self._params), ctypes.byref(self._buffers), self._stream()))
            self._stats_pending = False

    @property
    def stats(self) -> torch.Tensor:
        """Statistics of the last post step ([HA_STATS_LEN] float32, ``HA_HAND_STAT_*``)."""
        self.finalize_stats()
        return self._stats

    @property
    def consecutive_successes(self) -> torch.Tensor:
        """Running average of the successes of finished episodes ([1] float32, shadow_hand.py:795-798)."""
        self.finalize_stats()
        return self._consecutive_successes

    def compute_observations(self):
        """compute_observations only (VecTask.reset())."""
        self._require_engine()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_hand_observe(ctypes.byref(self._observe_params), ctypes.byref(self._buffers), self._stream()))

    def reset_idx(self, env_ids: torch.Tensor, goal_env_ids: Optional[torch.Tensor] = None) -> None:
        """shadow_hand.py:604-668 for an explicit id list (reset_done path)."""
        self._require_engine()
        saved = self.reset_buf.clone()
        self.reset_buf.zero_()
        self.reset_buf[env_ids] = 1
        self._rebind_sim_state()
        _lib.check(self._lib.ha_hand_reset_idx(ctypes.byref(self._params), ctypes.byref(self._buffers), self._rand_ptr(), self._stream()))
        self._stats_pending = False
        saved[env_ids] = 0
        self.reset_buf.copy_(saved)

    def get_state(self):
        # states_buf was clipped by the kernel (vec_task.py:374-376)
        return self.states_buf if self.rl_device == self.device else self.states_buf.to(self.rl_device)

    def bytes_per_env_step(self) -> int:
        return int(self._lib.ha_hand_bytes_per_env_step(ctypes.byref(self._params)))


class ShadowHand(_HandTask):
    """shadow_hand.py:42."""
    is_shadow = True
    num_obs_dict = {"openai": 42, "full_no_vel": 77, "full": 157, "full_state": 211}
    default_layout = SHADOW_HAND_LAYOUT
    dof_lower, dof_upper = SHADOW_HAND_LOWER, SHADOW_HAND_UPPER
    actuated = SHADOW_HAND_ACTUATED
    object_start = (0.0, -0.39, 0.54)


class AllegroHand(_HandTask):
    """allegro_hand.py:41."""
    is_shadow = False
    num_obs_dict = {"full_no_vel": 50, "full": 72, "full_state": 88}
    default_layout = ALLEGRO_HAND_LAYOUT
    dof_lower, dof_upper = ALLEGRO_HAND_LOWER, ALLEGRO_HAND_UPPER
    actuated = list(range(16))
    object_start = (0.0, -0.19, 0.56)      # hand pose (0, 0, 0.5) + startObjectPoseDY / DZ (allegro_hand.py:283-292)
