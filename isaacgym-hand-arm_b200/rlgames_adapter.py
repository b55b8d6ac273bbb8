"""rl_games-side adapter for dict-observation tasks (``ComplexObsRLGPUEnv``, ``isaacgymenvs/utils/rlgames_utils.py:313-412``).

Same surface -- ``step`` / ``reset`` return ``{"obs": ..., "states": ...}`` built from ``obs_spec`` (``train.py:162-167``),
``get_env_info`` reports the spaces -- but for specs with ``concat: True`` the concatenated tensors are not built with ``torch.cat``
every step: the task's post-physics kernel writes them directly (``AllegroHandDextremeADR.set_obs_views``)."""
from __future__ import annotations

from typing import Any, Dict, Tuple

import numpy as np
import torch

from .vec_task import Box


class ComplexObsRLGPUEnv:
    def __init__(self, env, obs_spec: Dict[str, Dict]):
        self.env, self.obs_spec = env, obs_spec
        if hasattr(env, "set_obs_views") and torch.device(env.device).type == "cuda":
            env.set_obs_views(obs_spec)

    def gen_obs_dict(self, key: str, obs_dict: Dict[str, torch.Tensor], obs_names, concat: bool):
        if not concat:
            return {k: obs_dict[k] for k in obs_names}
        views = getattr(self.env, "obs_views", {})
        if key in views:
            return views[key].to(self.env.rl_device)                  # written by the kernel: no per-step concatenation
        return torch.cat([obs_dict[name] for name in obs_names], dim=1)

    def _generate_obs(self, env_obs: Dict[str, torch.Tensor]) -> Dict[str, Any]:
        return {k: self.gen_obs_dict(k, env_obs, v["names"], v["concat"]) for k, v in self.obs_spec.items()}

    def step(self, action: torch.Tensor) -> Tuple[Dict[str, Any], torch.Tensor, torch.Tensor, Dict[str, Any]]:
        env_obs, rewards, dones, infos = self.env.step(action)
        return self._generate_obs(env_obs), rewards, dones, infos

    def reset(self) -> Dict[str, Any]:
        return self._generate_obs(self.env.reset())

    def get_number_of_agents(self) -> int:
        return getattr(self.env, "num_agents", 1)

    def gen_obs_space(self, obs_names, concat: bool):
        dims = self.env.cfg["env"]["obsDims"]
        if concat:
            return Box(low=-np.inf, high=np.inf, shape=(sum(dims[n][0] for n in obs_names),), dtype=np.float32)
        return {k: Box(low=-np.inf, high=np.inf, shape=dims[k], dtype=np.float32) for k in obs_names}

    def get_env_info(self) -> Dict[str, Any]:
        info = {"action_space": self.env.action_space}
        for v in self.obs_spec.values():
            info[v["space_name"]] = self.gen_obs_space(v["names"], v["concat"])
        return info
