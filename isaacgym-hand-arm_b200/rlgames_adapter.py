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


class RLGPUAlgoObserver:
    """``RLGPUAlgoObserver`` (``rlgames_utils.py:129-218``) with the ``episode_cumulative`` bookkeeping on the device.

    Same hooks and attributes (``process_infos(infos, done_indices)``, ``after_print_stats(frame, epoch_num, total_time)``,
    ``direct_info``, ``ep_infos``, ``new_finished_episodes``).  The reference adds every term to a running sum with one launch per
    term and then walks the finished envs in Python with one ``.item()`` per env and term (``:161-172``); here one kernel per step
    (``ha_observer_step``) does both from the done MASK, the deques are device ring buffers of ``games_to_track`` entries, and their
    mean / min / max (``:199-201``) come from ``ha_observer_summary`` with a single device->host copy at logging time.

    ``done_indices`` may be the index tensor rl_games passes (``dones.nonzero()``) or ``None`` when ``infos["dones"]`` carries
    the mask itself (no ``nonzero`` sync at all).  The terms are read in place when the dict values are the rows of one
    ``[n_terms, n_envs]`` block (as ``AllegroKukaBase.episode_cumulative`` is), else they are stacked first."""

    def __init__(self, games_to_track: int = 100):
        self.algo = None
        self.writer = None
        self.games_to_track = games_to_track
        self.ep_infos = []
        self.direct_info = {}
        self.new_finished_episodes = False
        self._keys = None
        self._lib = None

    def after_init(self, algo):
        self.algo = algo
        self.writer = algo.writer
        tracked = getattr(algo, "games_to_track", self.games_to_track)
        if self._keys is not None and tracked != self.games_to_track:
            # the device rings were sized by an earlier process_infos: a different window needs new rings (the deques of the
            # reference are created lazily with algo.games_to_track, :168)
            self._keys = None
            self._logged = 0
        self.games_to_track = tracked

    # ---- device state ---------------------------------------------------------------------------------------------------
    def _setup(self, terms: Dict[str, torch.Tensor]):
        import ctypes
        from . import _lib
        first = next(iter(terms.values()))
        if first.device.type != "cuda":
            raise _lib.HandArmLibraryError("RLGPUAlgoObserver needs CUDA tensors: there is no CPU fallback")
        self._lib = _lib.load()
        self._keys = list(terms.keys())
        K, N, G, dev = len(self._keys), first.shape[0], self.games_to_track, first.device
        self._n, self._k = N, K
        self.cumulative = torch.zeros((K, N), dtype=torch.float32, device=dev)
        self.episode_cumulative = {k: self.cumulative[i] for i, k in enumerate(self._keys)}        # reference attribute, row views
        self.ring = torch.zeros((K, G), dtype=torch.float32, device=dev)
        self._state = torch.zeros(4, dtype=torch.int64, device=dev)
        self._block_counts = torch.zeros(_lib.CONSTS["HA_OBSERVER_MAX_CTAS"], dtype=torch.int64, device=dev)
        self._summary = torch.zeros((3, K), dtype=torch.float32, device=dev)
        self._mask = torch.zeros(N, dtype=torch.int64, device=dev)
        b = _lib.HaObserverBuffers()
        b.cumulative, b.ring, b.state = self.cumulative.data_ptr(), self.ring.data_ptr(), self._state.data_ptr()
        b.block_counts, b.summary = self._block_counts.data_ptr(), self._summary.data_ptr()
        self._b, self._ct = b, ctypes

    def _terms_block(self, terms: Dict[str, torch.Tensor]) -> torch.Tensor:
        vals = [terms[k] for k in self._keys]
        base, step = vals[0].data_ptr(), self._n * 4
        if all(v.dtype == torch.float32 and v.is_contiguous() and v.data_ptr() == base + i * step for i, v in enumerate(vals)):
            return vals[0]                                                # rows of one [K, N] block: read in place
        self._stacked = torch.stack([v.float() for v in vals])
        return self._stacked

    def process_infos(self, infos, done_indices=None):
        assert isinstance(infos, dict), "RLGPUAlgoObserver expects dict info"
        from . import _lib
        if "episode" in infos:
            self.ep_infos.append(infos["episode"])
        if "episode_cumulative" in infos:
            terms = infos["episode_cumulative"]
            if self._keys is None:
                self._setup(terms)
            block = self._terms_block(terms)
            if done_indices is None:
                mask = infos["dones"]
                mask = mask if mask.dtype == torch.int64 else mask.to(torch.int64)
            else:
                mask = self._mask
                mask.zero_()
                mask[done_indices.reshape(-1)] = 1
            self._b.terms, self._b.dones = block.data_ptr(), mask.data_ptr()
            stream = self._ct.c_void_p(torch.cuda.current_stream(block.device).cuda_stream)
            _lib.check(self._lib.ha_observer_step(self._n, self._k, self.games_to_track, self._ct.byref(self._b), stream))
        # nested infos -> summary keys, scalars only (:175-182)
        if len(infos) > 0:
            self.direct_info = {}
            for k, v in _flatten_dict(infos).items():
                if isinstance(v, (float, int)) or (isinstance(v, torch.Tensor) and v.dim() == 0):
                    self.direct_info[k] = v

    def episodes_finished(self) -> int:
        return int(self._state[0].item()) if self._keys is not None else 0

    def episode_cumulative_summary(self) -> Dict[str, Tuple[float, float, float]]:
        """{key: (mean, min, max)} over the last ``games_to_track`` finished episodes (``:199-201``); one device->host copy."""
        from . import _lib
        if self._keys is None:
            return {}
        stream = self._ct.c_void_p(torch.cuda.current_stream(self.ring.device).cuda_stream)
        _lib.check(self._lib.ha_observer_summary(self._k, self.games_to_track, self._ct.byref(self._b), stream))
        s = self._summary.cpu()
        return {k: (float(s[0, i]), float(s[1, i]), float(s[2, i])) for i, k in enumerate(self._keys)}

    def after_print_stats(self, frame, epoch_num, total_time):
        if self.ep_infos:
            for key in self.ep_infos[0]:
                vals = [torch.as_tensor(e[key], dtype=torch.float32).reshape(-1).cpu() for e in self.ep_infos]
                self.writer.add_scalar("Episode/" + key, torch.mean(torch.cat(vals)), epoch_num)
            self.ep_infos.clear()
        appended = self.episodes_finished()
        if appended > getattr(self, "_logged", 0):            # the reference sets the flag per finished env in process_infos (:164);
            self.new_finished_episodes = True                 # here the finished envs are counted on the device and read once per print
            self._logged = appended
        if self.new_finished_episodes:                        # log these if and only if we have new finished episodes (:196-202)
            for key, (mean, mn, mx) in self.episode_cumulative_summary().items():
                self.writer.add_scalar(f"episode_cumulative/{key}", mean, frame)
                self.writer.add_scalar(f"episode_cumulative_min/{key}_min", mn, frame)
                self.writer.add_scalar(f"episode_cumulative_max/{key}_max", mx, frame)
            self.new_finished_episodes = False
        for k, v in self.direct_info.items():
            self.writer.add_scalar(f"{k}/frame", v, frame)
            self.writer.add_scalar(f"{k}/iter", v, epoch_num)
            self.writer.add_scalar(f"{k}/time", v, total_time)
        # custom data of the env (:206-214)
        env = getattr(getattr(self.algo, "vec_env", None), "env", None)
        if env is not None and hasattr(env, "log_data"):
            for k, v in env.log_data.items():
                self.writer.add_scalar(k + "/frame", v, frame)
                self.writer.add_scalar(k + "/iter", v, epoch_num)
                self.writer.add_scalar(k + "/time", v, total_time)


def _flatten_dict(d, prefix="", separator="/"):
    """``isaacgymenvs/utils/utils.py`` flatten_dict: nested dict -> {"a/b": v}."""
    out = {}
    for k, v in d.items():
        if isinstance(v, dict):
            out.update(_flatten_dict(v, prefix + k + separator, separator))
        else:
            out[prefix + k] = v
    return out
