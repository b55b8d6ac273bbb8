"""AllegroHandDextremeADR on the fused sm_100a kernels.

Drop-in for the per-step tensor path of ``isaacgymenvs/tasks/dextreme/allegro_hand_dextreme.py`` (``AllegroHandDextremeADR``,
``:1347``) behind the dict-observation contract of ``adr_vec_task.py:259-333``: same constructor, config keys, attribute names and
``extras`` keys; ``step`` returns a dict of the 20 named observation tensors (``dict_obs_cls = True``).  ``pre_physics_step`` and
``post_physics_step`` are one kernel launch each (``ha_dex_pre_step`` / ``ha_dex_post_step``); nothing on the step path
synchronises with the host.

Layout choices (B200-first):
* the 20 dict entries are column bands of ONE ``[N, 285]`` row buffer (``obs_rows``), written already clipped to
  ``+-clipObservations`` -- ``step`` returns views, the per-entry ``clone(clamp(...))`` of ``adr_vec_task.py:317-322`` disappears;
* the action queue (``prev_actions_queue`` ``[N, L, 16]`` shifted every step, ``:969``) is a slot-major ring buffer
  ``[L, N, 16]``; ``prev_actions_queue`` is offered as a read-only reconstruction for callers that want the reference view.

ADR: the per-env half of ``apply_randomizations`` -- ``recycle_envs``, ``sample_adr_tensor`` for every tensorised parameter and the
Dextreme draws (action latency, cube-pose refresh rate / offset, six affine tensors; ``:1392-1416``) -- is one more kernel
(``ha_dex_adr_resample``), launched ahead of K_pre on the envs in ``reset_buf & ~apply_reset_buf`` without reading the mask back.
The range update (``adr_update``: host deques and thresholds) and the Isaac Gym property randomisation are host control logic
outside the path: ``adr_params[name]["range"]`` is plain host state a caller updates, ``refresh_adr_table()`` pushes it.
"""
from __future__ import annotations

import ctypes
import math
from typing import Dict, Optional

import torch

from .. import _lib
from ..sim import ALLEGRO_HAND_LOWER, ALLEGRO_HAND_UPPER, ActorLayout, SyntheticSim
from ..vec_task import VecTask

C = _lib.CONSTS
DEXTREME_LAYOUT = ActorLayout(actors_per_env=3, bodies_per_env=19, dofs_per_env=16, hand_actor=(0,), object_actor=1, goal_actor=2,
                              palm_body=(0,), tip_body=(4, 8, 12, 16), object_body=17)
REWARD_KEYS = ["dist_rew", "rot_rew", "action_penalty", "action_delta_penalty", "velocity_penalty", "reach_goal_rew", "fall_rew", "timeout_rew"]
ADR_HOT_KEYS = ["action_delay_prob", "affine_action_white", "affine_cube_pose_white", "cube_obs_delay_prob", "affine_dof_pos_white", "rna_alpha"]
AFFINE_ADR_KEYS = ["affine_action_scaling", "affine_action_additive", "affine_cube_pose_scaling", "affine_cube_pose_additive",
                   "affine_dof_pos_scaling", "affine_dof_pos_additive"]


class AllegroHandDextremeADR(VecTask):
    dict_obs_cls = True
    default_layout = DEXTREME_LAYOUT

    def __init__(self, cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture=False, force_render=False,
                 sim: Optional[SyntheticSim] = None, layout: Optional[ActorLayout] = None, seed: int = 1234, snapshots: int = 1):
        self.cfg = cfg
        self._injected_sim, self._layout, self._seed, self._snapshots = sim, layout or self.default_layout, seed, snapshots
        self._read_cfg()
        self.fingertips = [s + "_link_3" for s in ["index", "middle", "ring", "thumb"]]
        self.num_fingertips = len(self.fingertips)
        self.num_obs_dict = self.get_num_obs_dict(16)
        cfg["env"]["obsDims"] = {k: (w,) for k, w in self.num_obs_dict.items()}
        self.up_axis = "z"
        self.asymmetric_obs = cfg["env"]["asymmetric_observations"]
        cfg["env"]["numActions"] = 16
        cfg["env"]["numObservations"] = sum(self.num_obs_dict.values())     # the row buffer; the dict entries are its column bands
        cfg["env"]["numStates"] = 0
        adr = cfg["task"].get("adr", {})
        self.use_adr = bool(adr.get("use_adr", False))
        self.adr_params = adr.get("params", {})
        self.worker_adr_boundary_fraction = adr.get("worker_adr_boundary_fraction", 0.4)
        self.adr_extended_boundary_sample = bool(adr.get("adr_extended_boundary_sample", False))
        self.update_adr_ranges = bool(adr.get("update_adr_ranges", False))       # adr_vec_task.py:501-515
        self.adr_queue_threshold_length = adr.get("adr_queue_threshold_length", 256)
        self.adr_objective_threshold_low = adr.get("adr_objective_threshold_low", 5)
        self.adr_objective_threshold_high = adr.get("adr_objective_threshold_high", 20)
        self.adr_clear_other_queues = bool(adr.get("clear_other_queues", False))
        self.adr_update_on_step = True       # run adr_update (one device->host copy per call) ahead of the resampling kernel
        self._adr_updater = None
        for k, prm in self.adr_params.items():                               # adr_vec_task.py:530-545
            prm["range"] = list(prm["init_range"])
            prm.setdefault("limits", [None, None])
            if prm.get("type", "uniform") != "uniform":
                raise NotImplementedError(f"ADR parameter {k}: only uniform tensorised parameters are built")
        self.adr_params_keys = list(self.adr_params.keys())
        self.adr_tensor_keys = [k for k in self.adr_params_keys if "range_path" not in self.adr_params[k]]
        self.num_adr_params = len(self.adr_params)
        self.first_randomization = True
        super().__init__(cfg, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture, force_render)
        self.dt = self.sim_params_dt
        if self.reset_time > 0.0:
            self.max_episode_length = int(round(self.reset_time / (self.control_freq_inv * self.dt)))
        self._allocate_task_state()
        self._build_abi()

    # ---- configuration (allegro_hand_dextreme.py:1083-1200, :1491-1497) --------------------------------
    def _read_cfg(self):
        env = self.cfg["env"]
        self.randomize = self.cfg["task"]["randomize"]
        self.randomization_params = self.cfg["task"]["randomization_params"]
        self.aggregate_mode = env["aggregateMode"]
        self.dist_reward_scale = env["distRewardScale"]
        self.rot_reward_scale = env["rotRewardScale"]
        self.action_penalty_scale = env["actionPenaltyScale"]
        self.action_delta_penalty_scale = env["actionDeltaPenaltyScale"]
        self.success_tolerance = env["successTolerance"]
        self.reach_goal_bonus = env["reachGoalBonus"]
        self.fall_dist = env["fallDistance"]
        self.fall_penalty = env["fallPenalty"]
        self.rot_eps = env["rotEps"]
        self.vel_obs_scale = 1.0            # the ADR subclass overrides the base class's 0.2 / 10.0 (:1494-1495)
        self.force_torque_obs_scale = 1.0
        self.max_effort = env.get("max_effort", 0.35)
        self.reset_position_noise = env["resetPositionNoise"]
        self.reset_position_noise_z = env["resetPositionNoiseZ"]
        self.reset_rotation_noise = env["resetRotationNoise"]
        self.reset_dof_pos_noise = env["resetDofPosRandomInterval"]
        self.reset_dof_vel_noise = env["resetDofVelRandomInterval"]
        self.force_scale = env.get("forceScale", 0.0)
        self.force_prob_range = env.get("forceProbRange", [0.001, 0.1])
        self.force_decay = env.get("forceDecay", 0.99)
        self.force_decay_interval = env.get("forceDecayInterval", 0.08)
        self.dof_speed_scale = env["dofSpeedScale"]
        self.use_relative_control = env["useRelativeControl"]
        if self.use_relative_control:
            raise NotImplementedError("useRelativeControl: the reference branch (:978-981) reads an undefined attribute and cannot run")
        self.use_capped_dof_control = env["use_capped_dof_control"]
        self.max_dof_radians_per_second = env["max_dof_radians_per_second"]
        self.num_success_hold_steps = env.get("num_success_hold_steps", 1)
        ama = env["actionsMovingAverage"]
        self.act_moving_average_range = ama["range"]
        self.act_moving_average_scheduled_steps = ama["schedule_steps"]
        self.act_moving_average_scheduled_freq = ama["schedule_freq"]
        self.act_moving_average_lower, self.act_moving_average_upper = self.act_moving_average_range
        self.act_moving_average = self.act_moving_average_upper
        rco = env.get("random_cube_observation")
        self.enable_random_obs = bool(rco and rco["enable"])
        self.random_cube_pose_prob = rco["prob"] if rco else 0.0
        self.apply_random_quat = env.get("apply_random_quat", True)
        self.debug_viz = False
        self.max_episode_length = env["episodeLength"]
        self.reset_time = env.get("resetTime", -1.0)
        self.print_success_stat = env["printNumSuccesses"]
        self.max_consecutive_successes = env["maxConsecutiveSuccesses"]
        self.av_factor = env.get("averFactor", 0.1)
        self.action_delay_prob_max = env["actionDelayProbMax"]
        self.action_latency_max = env["actionLatencyMax"]
        self.action_latency_scheduled_steps = env["actionLatencyScheduledSteps"]
        self.frame_start = 0
        self.object_type = env["objectType"]
        assert self.object_type in ["block", "egg"]
        rna = env.get("random_network_adversary")
        self.enable_rna = bool(rna and rna["enable"])
        self.control_freq_inv_cfg = env.get("controlFrequencyInv", 1)

    def get_num_obs_dict(self, num_dofs):
        """:168-199 -- insertion order is the column order of ``obs_rows`` (``HA_DEX_OBS_*``)."""
        return {"dof_pos": num_dofs, "dof_pos_randomized": num_dofs, "dof_vel": num_dofs, "dof_force": num_dofs, "object_vels": 6,
                "last_actions": num_dofs, "cube_random_params": 3, "hand_random_params": 1, "gravity_vec": 3,
                "ft_states": 13 * self.num_fingertips, "ft_force_torques": 6 * self.num_fingertips, "rb_forces": 3, "rot_dist": 2,
                "stochastic_delay_params": 4, "affine_params": 16 * 2 + 7 * 2 + 16 * 2, "object_pose": 7, "goal_pose": 7,
                "goal_relative_rot": 4, "object_pose_cam_randomized": 7, "goal_relative_rot_cam_randomized": 4}

    # ---- construction ------------------------------------------------------------------------
    def create_sim(self):
        """Replaces asset loading / actor creation (:201-436) by the synthetic producer."""
        L = self._layout
        self.up_axis_idx = 2
        self.sim = self._injected_sim or SyntheticSim(self.num_envs, L, self.device, seed=self._seed, snapshots=self._snapshots)
        self.gym = self.sim
        N, dev = self.num_envs, self.device
        self.num_hand_dofs = 16
        self.num_bodies, self.num_dofs = L.bodies_per_env, L.dofs_per_env
        self.hand_dof_lower_limits = torch.tensor(ALLEGRO_HAND_LOWER, dtype=torch.float, device=dev)
        self.hand_dof_upper_limits = torch.tensor(ALLEGRO_HAND_UPPER, dtype=torch.float, device=dev)
        self.hand_dof_default_pos = torch.zeros(16, dtype=torch.float, device=dev)
        self.hand_dof_default_vel = torch.zeros(16, dtype=torch.float, device=dev)
        self.actuated_dof_indices = torch.arange(16, dtype=torch.long, device=dev)
        self.fingertip_handles = torch.tensor(list(L.tip_body), dtype=torch.long, device=dev)
        self.palm_link_handle = (L.palm_body[0],)
        self.object_rb_handles = torch.tensor([L.object_body], dtype=torch.long, device=dev)
        self.object_rb_masses = torch.tensor([0.05], dtype=torch.float, device=dev)
        ar = torch.arange(N, dtype=torch.long, device=dev) * L.actors_per_env
        self.hand_indices, self.object_indices, self.goal_object_indices = ar + L.hand_actor[0], ar + L.object_actor, ar + L.goal_actor
        env = self.cfg["env"]
        init = torch.zeros(N, 13, dtype=torch.float)
        init[:, 0:3] = torch.tensor([0.0, env["startObjectPoseDY"], 0.5 + env["startObjectPoseDZ"]])     # hand pose (0, 0, 0.5) + offsets (:297-308)
        init[:, 6] = 1.0
        self.object_init_state = init.to(dev)
        self.goal_displacement_tensor = torch.tensor([-0.2, -0.06, 0.12], dtype=torch.float, device=dev)
        self.goal_states = self.object_init_state.clone()
        self.goal_states[:, self.up_axis_idx] -= 0.04
        self.goal_init_state = self.goal_states.clone()
        self.sim.fill_hand(init[0, 0:3].tolist(), self.goal_states, ALLEGRO_HAND_LOWER, ALLEGRO_HAND_UPPER)

    def _adr_init(self, name: str, gen: torch.Generator) -> torch.Tensor:
        lo, hi = (float(v) for v in self.adr_params.get(name, {}).get("init_range", [0.0, 0.0]))
        return (lo + (hi - lo) * torch.rand(self.num_envs, generator=gen)).to(self.device)

    def _allocate_task_state(self):
        """:1202-1329 and :1349-1370, same names and dtypes (the action queue excepted, see the module docstring)."""
        N, dev = self.num_envs, self.device
        f = dict(dtype=torch.float, device=dev)
        i64 = dict(dtype=torch.long, device=dev)
        gen = torch.Generator().manual_seed(self._seed)
        self.cube_random_params = torch.zeros((N, 3), **f)
        self.hand_random_params = torch.zeros((N, 1), **f)
        self.gravity_vec = torch.zeros((N, 3), **f)
        self.cube_pose_refresh_rate = torch.ones(N, **i64)          # sampled >= 1 by apply_randomizations before the first step (:1418-1427)
        self.cube_pose_refresh_offset = torch.zeros(N, **i64)
        self.action_latency = torch.zeros(N, **i64)
        self.affine_actions_scaling = torch.ones(N, 16, **f)
        self.affine_actions_additive = torch.zeros(N, 16, **f)
        self.affine_cube_pose_scaling = torch.ones(N, 7, **f)
        self.affine_cube_pose_additive = torch.zeros(N, 7, **f)
        self.affine_dof_pos_scaling = torch.ones(N, 16, **f)
        self.affine_dof_pos_additive = torch.zeros(N, 16, **f)
        keys = list(dict.fromkeys(self.adr_tensor_keys + ADR_HOT_KEYS))
        self.adr_tensor_values: Dict[str, torch.Tensor] = {k: self._adr_init(k, gen) for k in keys}
        self.worker_types = torch.zeros(N, **i64)
        self.adr_modes = torch.zeros(N, **i64)
        self.ext_rand_adr: Optional[torch.Tensor] = None               # parity mode: [N, HA_DEX_ADR_RND_STRIDE] raw draws
        self.all_env_ids = torch.arange(N, **i64)
        self.prev_targets = torch.zeros((N, 16), **f)
        self.cur_targets = torch.zeros((N, 16), **f)
        self.prev_prev_targets = torch.zeros((N, 16), **f)
        self.actions = torch.zeros((N, 16), **f)
        self.prev_actions = torch.zeros((N, 16), **f)
        self.queue_len = int(self.action_latency_max) + 1
        self.action_queue = torch.zeros((self.queue_len, N, 16), **f)
        self.dof_delta = torch.zeros((N, 16), **f)
        self.reset_goal_buf = self.reset_buf.clone()
        self.hold_count_buf = self.progress_buf.clone()
        self.apply_reset_buf = torch.zeros(N, **i64)
        self.successes = torch.zeros(N, **f)
        self.consecutive_successes = torch.zeros(1, **f)
        self.best_rotation_dist = -torch.ones(N, **f)
        self.total_successes = self.total_resets = 0
        self.force_decay_t = torch.tensor(self.force_decay, **f)
        self.force_prob_range_t = torch.tensor(self.force_prob_range, **f)
        u0 = torch.rand(N, generator=gen).to(dev)
        self.random_force_prob = torch.exp((torch.log(self.force_prob_range_t[0]) - torch.log(self.force_prob_range_t[1])) * u0
                                           + torch.log(self.force_prob_range_t[1]))
        self.rb_forces = torch.zeros((N, self.num_bodies, 3), **f)
        self.object_pose = self.sim.root_rows()[:, self._layout.object_actor, 0:7].clone().contiguous()
        self.obs_object_pose_freq = self.object_pose.clone()
        self.obs_object_pose = self.object_pose.clone()
        self._episode_cumulative = torch.zeros((len(REWARD_KEYS), N), **f)
        self.rewards_episode = {k: torch.zeros(N, **f) for k in REWARD_KEYS}
        self.obs_rows = self.obs_buf                                   # [N, 285]
        ofs = 0
        self.obs_dict_views: Dict[str, torch.Tensor] = {}
        for k, w in self.num_obs_dict.items():
            self.obs_dict_views[k] = self.obs_rows[:, ofs:ofs + w]
            ofs += w
        # engine-side buffers
        self.root_reset_indices = torch.zeros(2 * N, dtype=torch.int32, device=dev)
        self.hand_reset_indices = torch.zeros(N, dtype=torch.int32, device=dev)
        self.reset_counts = torch.zeros(4, dtype=torch.int32, device=dev)
        self.stats = torch.zeros(C["HA_STATS_LEN"], **f)
        self._stat_accum = torch.zeros(C["HA_STATS_LEN"], dtype=torch.int64, device=dev)
        self._tickets = torch.zeros(8, dtype=torch.int32, device=dev)
        self._object_actor_i32 = self.object_indices.to(torch.int32)
        self._goal_actor_i32 = self.goal_object_indices.to(torch.int32)
        self._hand_actor_i32 = self.hand_indices.to(torch.int32)
        self._dynamic = torch.zeros(C["HA_DEX_DYN_LEN"], **i64)
        self._dynamic[C["HA_DEX_DYN_RNG_SEED"]] = self._seed
        self.ext_rand: Optional[torch.Tensor] = None                   # parity mode: [N, HA_DEX_RND_STRIDE] raw draws (pre)
        self.ext_rand_post: Optional[torch.Tensor] = None              # parity mode: [N, HA_DEX_POST_RND_STRIDE] raw draws (post)
        self.rna_perturbation: Optional[torch.Tensor] = None           # [N, 16] DOF-space output of the RNA network
        # random network adversary (:417-430): 32 bins per joint between the limits, network on the device
        self.rna_network = None
        self.run_rna = True                                            # parity harnesses that feed rna_perturbation themselves clear this
        if self.enable_rna and torch.device(self.device).type == "cuda":
            from ..rna import RandomNetworkAdversary
            lo_l, hi_l = self.hand_dof_lower_limits.cpu(), self.hand_dof_upper_limits.cpu()
            self.discretised_dofs = torch.stack([torch.linspace(float(lo_l[i]), float(hi_l[i]), steps=32) for i in range(16)]).to(self.device)
            self.rna_network = RandomNetworkAdversary(self.num_envs, 16 + 7, 16, 32, self.device, seed=self._seed)
            self.random_adversary_weight_sample_freq = self.cfg["env"]["random_network_adversary"]["weight_sample_freq"]
        self.obs_views: Dict[str, torch.Tensor] = {}                   # concatenated views emitted by K_post (set_obs_views)
        self.adr_resample_on_step = True                               # launch the per-env ADR resampling ahead of every K_pre

    # ---- C-ABI structs -----------------------------------------------------------------------
    def _build_abi(self):
        if torch.device(self.device).type != "cuda":
            self._params = self._buffers = None
            return
        if self.sim.state_layout != "isaacgym":
            raise NotImplementedError("the DeXtreme kernels read Isaac Gym's AoS state tensors (HA_STATE_ISAACGYM)")
        self._lib = _lib.load()
        L = self._layout
        p = _lib.HaDexParams()
        p.abi_version = C["HA_ABI_VERSION"]
        p.n_envs, p.dofs_per_env, p.n_bodies = self.num_envs, self.num_dofs, self.num_bodies
        p.palm_body, p.object_body = L.palm_body[0], L.object_body
        for t, bd in enumerate(L.tip_body):
            p.tip_body[t] = bd
        p.max_episode_length = int(self.max_episode_length)
        p.max_consecutive_successes = int(self.max_consecutive_successes)
        p.num_success_hold_steps = int(self.num_success_hold_steps)
        p.queue_len = self.queue_len
        p.use_capped_dof_control = int(bool(self.use_capped_dof_control))
        p.apply_random_quat = int(bool(self.apply_random_quat))
        p.force_enabled = 1 if self.force_scale > 0.0 else 0
        ar = torch.arange(self.num_envs, dtype=torch.long, device=self.device) * L.actors_per_env
        regular = bool(torch.equal(self.object_indices, ar + L.object_actor)) and bool(torch.equal(self.goal_object_indices, ar + L.goal_actor))
        p.root_regular, p.actors_per_env, p.object_slot, p.goal_slot = int(regular), L.actors_per_env, L.object_actor, L.goal_actor
        lo, hi = self.hand_dof_lower_limits.cpu(), self.hand_dof_upper_limits.cpu()
        for d in range(16):
            p.dof_lower[d], p.dof_upper[d] = float(lo[d]), float(hi[d])
        self._lib.ha_dex_fill_derived(ctypes.byref(p))
        p.clip_obs = self.clip_obs if math.isfinite(self.clip_obs) else 0.0
        p.vel_obs_scale, p.force_torque_obs_scale = self.vel_obs_scale, self.force_torque_obs_scale
        p.dist_reward_scale, p.rot_reward_scale, p.rot_eps = self.dist_reward_scale, self.rot_reward_scale, self.rot_eps
        p.action_penalty_scale, p.action_delta_penalty_scale = self.action_penalty_scale, self.action_delta_penalty_scale
        p.success_tolerance, p.reach_goal_bonus = self.success_tolerance, self.reach_goal_bonus
        p.fall_dist, p.fall_penalty, p.av_factor = self.fall_dist, self.fall_penalty, self.av_factor
        p.max_dof_delta = self.max_dof_radians_per_second * self.dt * self.control_freq_inv
        p.force_decay_pow = float(torch.pow(self.force_decay_t.cpu(), self.dt / self.force_decay_interval))
        p.force_gain[0], p.force_gain[1] = float(self.object_rb_masses[0]), self.force_scale
        lr = self.force_prob_range_t.cpu()
        p.force_log_c0 = float(torch.log(lr[0]) - torch.log(lr[1]))
        p.force_log_c1 = float(torch.log(lr[1]))
        p.reset_position_noise, p.reset_position_noise_z = self.reset_position_noise, self.reset_position_noise_z
        p.reset_dof_pos_noise, p.reset_dof_vel_noise = self.reset_dof_pos_noise, self.reset_dof_vel_noise
        gd = self.goal_displacement_tensor.cpu()
        for c in range(3):
            p.goal_displacement[c] = float(gd[c])
        p.random_cube_pose_prob = self.random_cube_pose_prob
        self._params = p
        self._set_moving_average(self.act_moving_average)
        self._buffers = _lib.HaDexBuffers()
        self._bind_buffers()
        self._adr_table = self._adr_buffers = None
        if self.use_adr and all(k in self.adr_tensor_keys for k in ["action_latency", "cube_pose_refresh_rate"] + AFFINE_ADR_KEYS):
            ab = _lib.HaDexAdrBuffers()
            for i, k in enumerate(self.adr_tensor_keys):
                ab.values[i] = self.adr_tensor_values[k].data_ptr()
            ab.worker_types, ab.adr_modes = self.worker_types.data_ptr(), self.adr_modes.data_ptr()
            self._adr_buffers = ab
            self.refresh_adr_table()

    def refresh_adr_table(self):
        """Push the host-side ADR ranges (``adr_params[k]["range"]`` / ``["next_limits"]``) into the table the resampling kernel reads."""
        t = _lib.HaDexAdrTable()
        keys = self.adr_tensor_keys
        assert len(keys) <= C["HA_DEX_ADR_MAX_PARAMS"]
        t.n_params_total, t.n_tensor = self.num_adr_params, len(keys)
        for i, k in enumerate(keys):
            prm = self.adr_params[k]
            lo, hi = (float(v) for v in prm["range"])
            nlo, nhi = (float(v) for v in prm.get("next_limits") or (lo, hi))
            t.param_index[i] = self.adr_params_keys.index(k)
            t.lo[i], t.span[i] = lo, hi - lo
            t.low_value[i], t.high_value[i] = (nlo, nhi) if self.adr_extended_boundary_sample else (lo, hi)
        t.boundary_fraction = self.worker_adr_boundary_fraction
        t.idx_action_latency, t.idx_cube_pose_refresh_rate = keys.index("action_latency"), keys.index("cube_pose_refresh_rate")
        for a, k in enumerate(AFFINE_ADR_KEYS):
            t.idx_affine[a] = keys.index(k)
        self._adr_table = t

    def _set_moving_average(self, value: float):
        self.act_moving_average = value
        if self._params is not None:
            self._params.act_moving_average = value
            self._params.one_minus_act_moving_average = 1.0 - value

    def _bind_buffers(self):
        b, adr = self._buffers, self.adr_tensor_values
        self._bound = dict(
            root_state=self.sim.root_state, rigid_body_state=self.sim.rigid_body_state, dof_state=self.sim.dof_state,
            dof_force=self.sim.dof_force, force_sensors=self.sim.force_sensors, object_actor=self._object_actor_i32,
            goal_actor=self._goal_actor_i32, hand_actor=self._hand_actor_i32, object_init_state=self.object_init_state,
            goal_init_state=self.goal_init_state, action_latency=self.action_latency, cube_pose_refresh_rate=self.cube_pose_refresh_rate,
            cube_pose_refresh_offset=self.cube_pose_refresh_offset, adr_action_delay_prob=adr["action_delay_prob"],
            adr_affine_action_white=adr["affine_action_white"], adr_affine_cube_pose_white=adr["affine_cube_pose_white"],
            adr_cube_obs_delay_prob=adr["cube_obs_delay_prob"], adr_affine_dof_pos_white=adr["affine_dof_pos_white"],
            adr_rna_alpha=adr["rna_alpha"], affine_actions_scaling=self.affine_actions_scaling,
            affine_actions_additive=self.affine_actions_additive, affine_cube_pose_scaling=self.affine_cube_pose_scaling,
            affine_cube_pose_additive=self.affine_cube_pose_additive, affine_dof_pos_scaling=self.affine_dof_pos_scaling,
            affine_dof_pos_additive=self.affine_dof_pos_additive, cube_random_params=self.cube_random_params,
            hand_random_params=self.hand_random_params, gravity_vec=self.gravity_vec, goal_states=self.goal_states, obs_rows=self.obs_rows,
            rew_buf=self.rew_buf, reset_buf=self.reset_buf, reset_goal_buf=self.reset_goal_buf, progress_buf=self.progress_buf,
            hold_count_buf=self.hold_count_buf, apply_reset_buf=self.apply_reset_buf, timeout_buf=self.timeout_buf, successes=self.successes,
            consecutive_successes=self.consecutive_successes, best_rotation_dist=self.best_rotation_dist, prev_targets=self.prev_targets,
            cur_targets=self.cur_targets, prev_prev_targets=self.prev_prev_targets, actions=self.actions, prev_actions=self.prev_actions,
            action_queue=self.action_queue, dof_delta=self.dof_delta, rb_forces=self.rb_forces, random_force_prob=self.random_force_prob,
            object_pose=self.object_pose, obs_object_pose=self.obs_object_pose, obs_object_pose_freq=self.obs_object_pose_freq,
            episode_cumulative=self._episode_cumulative, root_reset_indices=self.root_reset_indices,
            hand_reset_indices=self.hand_reset_indices, reset_counts=self.reset_counts, stats=self.stats, stat_accum=self._stat_accum,
            tickets=self._tickets, dynamic=self._dynamic)
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
        self._dynamic[C["HA_DEX_DYN_RNG_SEED"]] = int(seed)
        self._dynamic[C["HA_DEX_DYN_RNG_STEP"]] = 0

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
        return self.sim.dof_state.view(self.num_envs, -1, 2)[:, :16]

    @property
    def dof_pos(self):
        return self.dof_state[..., 0]

    @property
    def dof_vel(self):
        return self.dof_state[..., 1]

    @property
    def dof_force_tensor(self):
        return self.sim.dof_force

    @property
    def vec_sensor_tensor(self):
        return self.sim.force_sensors

    @property
    def frame(self) -> int:
        """``self.frame`` of the reference (device counter; reading it synchronises)."""
        return int(self._dynamic[C["HA_DEX_DYN_FRAME"]])

    @property
    def object_pose_wrt_wrist(self):
        return self.obs_dict_views["object_pose"]

    @property
    def prev_actions_queue(self) -> torch.Tensor:
        """The reference's ``[N, L, 16]`` view of the ring buffer: entry k = the action of k steps ago (reads the head: synchronises)."""
        head = int(self._dynamic[C["HA_DEX_DYN_QUEUE_HEAD"]])
        order = [(head - k) % self.queue_len for k in range(self.queue_len)]
        return self.action_queue[order].permute(1, 0, 2).contiguous()

    def set_prev_actions_queue(self, queue: torch.Tensor) -> None:
        """Load a reference-layout ``[N, L, 16]`` queue into the ring buffer (tests, checkpoints)."""
        head = int(self._dynamic[C["HA_DEX_DYN_QUEUE_HEAD"]])
        for k in range(self.queue_len):
            self.action_queue[(head - k) % self.queue_len].copy_(queue[:, k].to(self.device))

    def get_adr_tensor(self, param_name, env_ids=None):
        v = self.adr_tensor_values[param_name]
        return v if env_ids is None else v[env_ids]

    def adr_update(self, adr_objective: torch.Tensor, all_envs: bool = False) -> None:
        """``ADRVecTask.adr_update`` (adr_vec_task.py:760-918) up to its recycle / resample tail (``ha_dex_adr_resample``, launched
        right after by ``apply_randomizations``): objective queues on the device, one copy to the host, scalar range logic there
        (``handarm_b200/adr.py``).  ``rand_envs`` is the mask the resampling kernel uses: ``reset_buf & ~apply_reset_buf``."""
        if self._adr_updater is None:
            from ..adr import AdrRangeUpdater
            self._adr_updater = AdrRangeUpdater(self.adr_params, self.num_envs, self.device, self.worker_types, self.adr_modes,
                                                self.adr_queue_threshold_length, self.adr_objective_threshold_low,
                                                self.adr_objective_threshold_high, self.adr_clear_other_queues)
        obj = adr_objective if adr_objective.dtype == torch.float32 else adr_objective.float()
        log = self.last_step % 100 == 0
        if self._adr_updater.update(self.reset_buf, self.apply_reset_buf, obj.contiguous(), all_envs=all_envs, extras=self.extras, log=log):
            self.refresh_adr_table()

    def get_rna_alpha(self):
        return self.get_adr_tensor("rna_alpha").view(-1, 1)

    def apply_randomizations(self, dr_params=None, randomize_buf=None, adr_objective=None, randomisation_callback=None):
        """The per-env tensor half of ADR ``apply_randomizations`` (adr_vec_task.py:740-758, :666-730 via :915-917;
        allegro_hand_dextreme.py:1392-1416) as one launch on the envs in ``reset_buf & ~apply_reset_buf`` (every env on the first
        call, like ``first_randomization``).  ``randomize_buf`` is accepted for signature parity; the kernel derives the same mask."""
        self.last_step = self.control_steps          # adr_vec_task.py:939: self.last_step = self.gym.get_frame_count(self.sim)
        if self._adr_table is None:
            return
        self._require_engine()
        if self.use_adr and self.update_adr_ranges and self.adr_update_on_step and adr_objective is not None:
            self.adr_update(adr_objective, all_envs=self.first_randomization)
        self._adr_table.all_envs = 1 if self.first_randomization else 0
        self.first_randomization = False
        if self.ext_rand_adr is not None:
            r = self.ext_rand_adr
            assert r.shape == (self.num_envs, C["HA_DEX_ADR_RND_STRIDE"]) and r.dtype == torch.float32 and r.is_contiguous()
            self._params.use_fast_rng, ptr = 0, ctypes.c_void_p(r.data_ptr())
        else:
            self._params.use_fast_rng, ptr = 1, ctypes.c_void_p(0)
        _lib.check(self._lib.ha_dex_adr_resample(ctypes.byref(self._params), ctypes.byref(self._buffers), ctypes.byref(self._adr_table),
                                                 ctypes.byref(self._adr_buffers), ptr, self._stream()))

    # ---- the hot path --------------------------------------------------------------------------
    def _rand_ptrs(self):
        if self.ext_rand is not None or self.ext_rand_post is not None:
            pre, post = self.ext_rand, self.ext_rand_post
            assert pre is not None and post is not None, "parity mode needs both ext_rand and ext_rand_post"
            assert pre.shape == (self.num_envs, C["HA_DEX_RND_STRIDE"]) and pre.dtype == torch.float32 and pre.is_contiguous()
            assert post.shape == (self.num_envs, C["HA_DEX_POST_RND_STRIDE"]) and post.dtype == torch.float32 and post.is_contiguous()
            self._params.use_fast_rng = 0
            return ctypes.c_void_p(pre.data_ptr()), ctypes.c_void_p(post.data_ptr())
        self._params.use_fast_rng = 1
        return ctypes.c_void_p(0), ctypes.c_void_p(0)

    def update_action_moving_average(self):
        """:896-909 -- host scalar schedule."""
        step = self.control_steps
        if step > 0 and step % self.act_moving_average_scheduled_freq == 0:
            sched = 1.0 / self.act_moving_average_scheduled_steps * min(step, self.act_moving_average_scheduled_steps)
            self._set_moving_average(self.act_moving_average_upper + (self.act_moving_average_lower - self.act_moving_average_upper) * sched)
            self.extras["annealing/action_moving_average_scalar"] = self.act_moving_average

    def pre_physics_step(self, actions: torch.Tensor):
        """:912-949 as one kernel."""
        self._require_engine()
        self.update_action_moving_average()
        if actions.dtype != torch.float32 or not actions.is_contiguous():
            actions = actions.float().contiguous()
        self._rebind_sim_state()
        if self.randomize and self.use_adr and self.adr_resample_on_step:
            self.apply_randomizations(self.randomization_params, None, adr_objective=self.successes)      # (:926-937), mask derived in-kernel
        if self.rna_network is not None and self.run_rna:                     # get_random_network_adversary_action (:864-876)
            if self.last_step > 0 and self.last_step % self.random_adversary_weight_sample_freq == 0:
                self.rna_network._refresh()
            self.rna_perturbation = self.rna_network.perturbation(self.sim.dof_state, self.object_pose_wrt_wrist, self.discretised_dofs)
        rna = self.rna_perturbation
        rna_ptr = ctypes.c_void_p(rna.data_ptr()) if rna is not None else ctypes.c_void_p(0)
        _lib.check(self._lib.ha_dex_pre_step(ctypes.byref(self._params), ctypes.byref(self._buffers), ctypes.c_void_p(actions.data_ptr()),
                                             rna_ptr, self._rand_ptrs()[0], self._stream()))
        sim = self.sim
        if sim.record_calls:
            sim.record("set_actor_root_state_tensor_indexed", sim.root_state, self.root_reset_indices, self.reset_counts[0:1])
            sim.record("set_dof_position_target_tensor_indexed", self.prev_targets, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_dof_state_tensor_indexed", sim.dof_state, self.hand_reset_indices, self.reset_counts[1:2])
            sim.record("set_dof_position_target_tensor", self.cur_targets)
            if self.force_scale > 0.0:
                sim.record("apply_rigid_body_force_tensors", self.rb_forces, None, "LOCAL_SPACE")

    def post_physics_step(self):
        """:1033-1048 + compute_reward bookkeeping (:438-464) + the step tail (adr_vec_task.py:289-325) as one kernel."""
        self._require_engine()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_dex_post_step(ctypes.byref(self._params), ctypes.byref(self._buffers), self._rand_ptrs()[1], self._stream()))
        self.extras["consecutive_successes"] = self.consecutive_successes[0]      # mean of the [1] tensor: a 0-dim view, no launch
        self.extras["true_objective"] = self.successes
        self.extras["episode_cumulative"] = {k: self._episode_cumulative[i] for i, k in enumerate(REWARD_KEYS)}

    def compute_observations(self):
        self._require_engine()
        self._rebind_sim_state()
        _lib.check(self._lib.ha_dex_observe(ctypes.byref(self._params), ctypes.byref(self._buffers), self._rand_ptrs()[1], self._stream()))

    def set_obs_views(self, obs_spec: Dict[str, Dict]) -> None:
        """Have K_post also emit the concatenated tensors the rl_games adapter asks for (``obs_spec`` as ``train.py:162-167`` builds it:
        ``{"obs": {"names": [...], "concat": True, ...}, "states": {...}}``), so ``ComplexObsRLGPUEnv._generate_obs``
        (``rlgames_utils.py:341-399``) needs no ``torch.cat``."""
        self._require_engine()
        ofs, o = {}, 0
        for k, w in self.num_obs_dict.items():
            ofs[k] = (o, w)
            o += w
        self.obs_views: Dict[str, torch.Tensor] = {}
        self._view_maps = []
        for v, key in enumerate(("obs", "states")):
            names = obs_spec.get(key, {}).get("names", []) if obs_spec.get(key, {}).get("concat", True) else []
            cols = [c for n in names for c in range(ofs[n][0], ofs[n][0] + ofs[n][1])]
            self._params.n_view_cols[v] = len(cols)
            if cols:
                m = torch.tensor(cols, dtype=torch.int32, device=self.device)
                rows = torch.zeros((self.num_envs, len(cols)), dtype=torch.float, device=self.device)
                self._view_maps.append(m)
                self.obs_views[key] = rows
                self._buffers.view_cols[v], self._buffers.view_rows[v] = m.data_ptr(), rows.data_ptr()

    def _obs_dict_out(self) -> Dict[str, torch.Tensor]:
        return {k: v.to(self.rl_device) for k, v in self.obs_dict_views.items()}

    def step(self, actions: torch.Tensor):
        """adr_vec_task.py:259-333 (dict observations)."""
        if actions.device != torch.device(self.device):
            actions = actions.to(self.device, non_blocking=True)
        action_tensor = actions if not math.isfinite(self.clip_actions) else torch.clamp(actions, -self.clip_actions, self.clip_actions)
        self.pre_physics_step(action_tensor)
        for _ in range(self.control_freq_inv):
            self.sim.simulate()
        self.post_physics_step()
        self.control_steps += 1
        self.extras["time_outs"] = self.timeout_buf.to(self.rl_device)
        return self._obs_dict_out(), self.rew_buf.to(self.rl_device), self.reset_buf.to(self.rl_device), self.extras

    def reset(self):
        """adr_vec_task.py:335-360: first observations only."""
        self.compute_observations()
        return self._obs_dict_out()
