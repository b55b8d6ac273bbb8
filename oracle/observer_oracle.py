"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the episode_cumulative half of ``RLGPUAlgoObserver``
(``isaacgymenvs/utils/rlgames_utils.py:149-172`` process_infos, ``:192-205`` after_print_stats).

Pinned bit-exactly against the reference's own class (``tests/test_observer_oracle_vs_reference.py``, which imports it
from ``/root/reference`` through ``oracle/reference_loader.load_rlgames_utils``).  Only ``tests/`` may import this module;
the product path is ``handarm_b200.rlgames_adapter.RLGPUAlgoObserver`` on top of ``ha_observer_step``.
"""
from collections import deque
from typing import Dict, List

import numpy as np
import torch


class ObserverOracle:
    def __init__(self, games_to_track: int):
        self.games_to_track = games_to_track
        self.episode_cumulative: Dict[str, torch.Tensor] = {}          # rlgames_utils.py:140
        self.episode_cumulative_avg: Dict[str, deque] = {}             # :141
        self.new_finished_episodes = False

    def process_infos(self, infos: Dict, done_indices: torch.Tensor) -> None:
        """:156-172."""
        if "episode_cumulative" not in infos:
            return
        for key, value in infos["episode_cumulative"].items():
            if key not in self.episode_cumulative:
                self.episode_cumulative[key] = torch.zeros_like(value)
            self.episode_cumulative[key] += value
        for done_idx in done_indices:
            self.new_finished_episodes = True
            i = int(done_idx.item())
            for key in infos["episode_cumulative"]:
                if key not in self.episode_cumulative_avg:
                    self.episode_cumulative_avg[key] = deque([], maxlen=self.games_to_track)
                self.episode_cumulative_avg[key].append(self.episode_cumulative[key][i].item())
                self.episode_cumulative[key][i] = 0

    def summary(self) -> Dict[str, List[float]]:
        """What after_print_stats logs per key (:199-201): [mean, min, max] of the deque."""
        return {k: [float(np.mean(d)), float(np.min(d)), float(np.max(d))] for k, d in self.episode_cumulative_avg.items()}


def make_case(n_envs: int, n_terms: int, steps: int, seed: int, done_prob: float = 0.02):
    """Seeded inputs: per step a dict of [n_envs] term tensors and a done mask (ascending done_indices as nonzero() gives them)."""
    g = torch.Generator().manual_seed(seed)
    keys = [f"term_{i}" for i in range(n_terms)]
    out = []
    for _ in range(steps):
        terms = {k: torch.randn(n_envs, generator=g) for k in keys}
        dones = (torch.rand(n_envs, generator=g) < done_prob).long()
        out.append((terms, dones))
    return keys, out
