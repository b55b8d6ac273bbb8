"""VecTask surface of the reference (``isaacgymenvs/tasks/base/vec_task.py``), kept as it is.

Same constructor signature, ``step`` / ``reset`` / ``reset_done`` contract, buffer names and dtypes
(``vec_task.py:329-354``: ``obs_buf`` f32, ``rew_buf`` f32, ``reset_buf`` i64 initialised to ones,
``progress_buf`` / ``randomize_buf`` i64), ``extras`` dict and the side channels
``get_env_state`` / ``set_env_state`` / ``set_train_info``.  What changes is underneath: the
simulator is the synthetic producer of ``sim.py`` and the two task hooks are one fused sm_100a
kernel each, so the ``step`` tail (``vec_task.py:424-432``: ``timeout_buf`` and the obs clip) is
folded into the post-physics kernel.
"""
from __future__ import annotations

import abc
from typing import Any, Dict, Optional, Tuple

import numpy as np
import torch


class Box:
    """Minimal stand-in for ``gym.spaces.Box`` (gym is not a dependency of the hot path)."""

    def __init__(self, low, high, shape=None, dtype=np.float32):
        self.shape = tuple(shape) if shape is not None else tuple(np.shape(low))
        self.low = np.broadcast_to(np.asarray(low, dtype=dtype), self.shape)
        self.high = np.broadcast_to(np.asarray(high, dtype=dtype), self.shape)
        self.dtype = np.dtype(dtype)

    def __repr__(self):
        return f"Box({self.low.min()}, {self.high.max()}, {self.shape}, {self.dtype})"


class Env(abc.ABC):
    """``vec_task.py:67-135``: counts, spaces and clip values parsed from the task config."""

    def __init__(self, config: Dict[str, Any], rl_device: str, sim_device: str, graphics_device_id: int, headless: bool):
        split_device = sim_device.split(":")
        self.device_type = split_device[0]
        self.device_id = int(split_device[1]) if len(split_device) > 1 else 0
        self.device = "cpu"
        if config["sim"]["use_gpu_pipeline"]:
            if self.device_type.lower() in ("cuda", "gpu"):
                self.device = "cuda:" + str(self.device_id)
            else:
                print("GPU Pipeline can only be used with GPU simulation. Forcing CPU Pipeline.")
                config["sim"]["use_gpu_pipeline"] = False
        self.rl_device = rl_device
        self.headless = headless
        self.graphics_device_id = -1 if headless else graphics_device_id   # no graphics stack: never used

        env = config["env"]
        self.num_environments = env["numEnvs"]
        self.num_agents = env.get("numAgents", 1)
        self.num_observations = env.get("numObservations", 0)
        self.num_teacher_observations = env.get("numTeacherObservations", 0)
        self.num_states = env.get("numStates", 0)
        self.num_actions = env["numActions"]
        self.control_freq_inv = env.get("controlFrequencyInv", 1)

        self.obs_space = Box(-np.inf, np.inf, (self.num_obs,))
        self.state_space = Box(-np.inf, np.inf, (self.num_states,))
        self.act_space = Box(-1.0, 1.0, (self.num_actions,))
        self.teacher_obs_space = Box(-np.inf, np.inf, (self.num_teacher_observations,))

        self.clip_obs = env.get("clipObservations", np.inf)
        self.clip_actions = env.get("clipActions", np.inf)

    # properties mirrored from vec_task.py:137-211
    @property
    def observation_space(self):
        return self.obs_space

    @property
    def action_space(self):
        return self.act_space

    @property
    def teacher_observation_space(self):
        return self.teacher_obs_space

    @property
    def num_envs(self) -> int:
        return self.num_environments

    @property
    def num_acts(self) -> int:
        return self.num_actions

    @property
    def num_obs(self) -> int:
        return self.num_observations

    @property
    def num_teacher_obs(self) -> int:
        return self.num_teacher_observations

    @property
    def observation_keys(self):
        # RLGPUEnv.get_env_info reads this unconditionally (rlgames_utils.py:276)
        return None

    @property
    def observation_start_end(self):
        return None

    @property
    def teacher_observation_start_end(self):
        return None

    def set_train_info(self, env_frames, *args, **kwargs):
        """vec_task.py:213-220: hook for adaptive curricula; default is a no-op."""

    def get_env_state(self):
        return None

    def set_env_state(self, env_state):
        pass


class VecTask(Env):
    """``vec_task.py:233-491`` with the simulator replaced by a tensor producer."""

    def __init__(self, config, rl_device, sim_device, graphics_device_id, headless, virtual_screen_capture: bool = False,
                 force_render: bool = False):
        super().__init__(config, rl_device, sim_device, graphics_device_id, headless)
        self.virtual_screen_capture = False   # graphics: unavailable on this box
        self.force_render = False
        self.viewer = None
        self.cfg = config
        self.sim_params_dt = float(config["sim"]["dt"])
        self.physics_engine = config.get("physics_engine", "physx")
        if self.physics_engine not in ("physx", "flex"):
            raise ValueError(f"Invalid physics engine backend: {self.physics_engine}")   # vec_task.py:264-265
        self.first_randomization = True
        self.original_props: Dict = {}
        self.dr_randomizations: Dict = {}
        self.actor_params_generator = None
        self.extern_actor_params: Dict = {}
        self.last_step = -1
        self.last_rand_step = -1
        self.control_steps = 0
        self._obs_noise_active = False
        self.return_obs_copy = bool(config["env"].get("returnObsCopy", False))
        self.create_sim()
        self.allocate_buffers()
        self.obs_dict: Dict[str, Any] = {}

    def allocate_buffers(self) -> None:
        """vec_task.py:329-354 (same names, shapes, dtypes, initial values)."""
        n, dev = self.num_envs, self.device
        self.obs_buf = torch.zeros((n, self.num_obs), device=dev, dtype=torch.float)
        self.teacher_obs_buf = torch.zeros((n, self.num_teacher_obs), device=dev, dtype=torch.float)
        self.states_buf = torch.zeros((n, self.num_states), device=dev, dtype=torch.float)
        self.rew_buf = torch.zeros(n, device=dev, dtype=torch.float)
        self.reset_buf = torch.ones(n, device=dev, dtype=torch.long)
        self.timeout_buf = torch.zeros(n, device=dev, dtype=torch.bool)
        self.progress_buf = torch.zeros(n, device=dev, dtype=torch.long)
        self.randomize_buf = torch.zeros(n, device=dev, dtype=torch.long)
        self.extras: Dict[str, Any] = {}

    @abc.abstractmethod
    def create_sim(self):
        ...

    @abc.abstractmethod
    def pre_physics_step(self, actions: torch.Tensor):
        ...

    @abc.abstractmethod
    def post_physics_step(self):
        ...

    def get_state(self):
        """vec_task.py:374-376."""
        return torch.clamp(self.states_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)

    def _obs_out(self) -> torch.Tensor:
        """The clip of vec_task.py:432 happened inside the post-physics kernel (or with the noise).  ``obs_dict["obs"]`` ALIASES
        ``obs_buf`` when rl_device == device: the next step overwrites it, where the reference hands out a fresh clamped tensor
        every call.  rl_games copies the observation into its experience buffer before stepping again, so the 15 MB copy per step
        is not paid by default; ``env.returnObsCopy: True`` (or ``task.return_obs_copy = True``) restores the reference's
        semantics for callers that keep the tensor across steps."""
        if self.rl_device != self.device:
            return self.obs_buf.to(self.rl_device)
        return self.obs_buf.clone() if self.return_obs_copy else self.obs_buf

    def _set_in_kernel_obs_clip(self, on: bool) -> None:
        """Task hook: fold the VecTask clip (vec_task.py:432) into the post kernel's store (``on``) or leave it to the caller
        (``off``: observation noise is active and the reference clips after the noise)."""

    def step(self, actions: torch.Tensor) -> Tuple[Dict[str, torch.Tensor], torch.Tensor, torch.Tensor, Dict[str, Any]]:
        """vec_task.py:390-441."""
        if self.dr_randomizations.get("actions", None):
            actions = self.dr_randomizations["actions"]["noise_lambda"](actions)
        if actions.device != torch.device(self.device):
            actions = actions.to(self.device, non_blocking=True)
        action_tensor = actions if not np.isfinite(self.clip_actions) else torch.clamp(actions, -self.clip_actions, self.clip_actions)
        obs_noise = self.dr_randomizations.get("observations", None)
        if bool(obs_noise) != self._obs_noise_active:      # the reference clips AFTER the noise: the in-kernel clip is off while it is on
            self._obs_noise_active = bool(obs_noise)
            self._set_in_kernel_obs_clip(not self._obs_noise_active)
        self.pre_physics_step(action_tensor)
        for _ in range(self.control_freq_inv):
            self.sim.simulate()
        self.post_physics_step()          # also fills timeout_buf and clips obs (vec_task.py:424, :432)
        finalize = getattr(self, "finalize_stats", None)
        if finalize is not None:          # extras of THIS step are exact when step() returns (a raw pre/post loop folds them one launch later)
            finalize()
        self.control_steps += 1
        if obs_noise:
            # vec_task.py:427-432: noise, then the clip.  Applied IN PLACE: obs_buf stays the tensor the kernels are bound to (the
            # reference rebinds self.obs_buf to the closure's result; rebinding here would leave the kernels writing elsewhere)
            clip = float(self.clip_obs) if np.isfinite(self.clip_obs) else 0.0
            obs_noise["noise_lambda"](self.obs_buf, out=self.obs_buf, clip=clip)
        self.extras["time_outs"] = self.timeout_buf.to(self.rl_device)
        self.obs_dict["obs"] = self._obs_out()
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, self.rew_buf.to(self.rl_device), self.reset_buf.to(self.rl_device), self.extras

    def apply_randomizations(self, dr_params):
        """The non-physical half of ``vec_task.py:646-754``: (re)builds the observation / action noise closures that ``step``
        applies.  Physics-property randomisation (``:756-876``) drives the Isaac Gym API and is outside the path."""
        from .dr_noise import build_noise_closures
        rand_freq = dr_params.get("frequency", 1)
        self.last_step = self.control_steps                      # gym.get_frame_count(sim)
        do_nonenv_randomize = self.first_randomization or (self.last_step - self.last_rand_step) >= rand_freq
        if do_nonenv_randomize:
            self.last_rand_step = self.last_step
            self.dr_randomizations.update(build_noise_closures(dr_params, self.last_step, seed=getattr(self, "_seed", 0)))
        self.first_randomization = False

    def zero_actions(self) -> torch.Tensor:
        return torch.zeros([self.num_envs, self.num_actions], dtype=torch.float32, device=self.rl_device)

    def reset_idx(self, env_idx):
        pass

    def compute_observations(self):
        pass

    def reset(self):
        """vec_task.py:459-474: first observations only; resets are deferred to the first step."""
        self.compute_observations()
        self.obs_dict["obs"] = torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict

    def reset_done(self):
        """vec_task.py:476-491."""
        done_env_ids = self.reset_buf.nonzero(as_tuple=False).flatten()
        if len(done_env_ids) > 0:
            self.reset_idx(done_env_ids)
        self.obs_dict["obs"] = torch.clamp(self.obs_buf, -self.clip_obs, self.clip_obs).to(self.rl_device)
        if self.num_states > 0:
            self.obs_dict["states"] = self.get_state()
        return self.obs_dict, done_env_ids

    def render(self, mode="rgb_array"):
        """Graphics are unavailable on this box (no GL/EGL/Vulkan loaders): headless only."""
        return None
