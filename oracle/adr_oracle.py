"""TEST INFRASTRUCTURE ONLY -- CPU restatement of ``ADRVecTask.adr_update`` (``isaacgymenvs/tasks/dextreme/adr_vec_task.py:760-918``)
up to (not including) the recycle / resample tail, with ``clear_other_queues: False`` (the shipped config).  Pinned against the
reference's own method in ``tests/test_adr_oracle_vs_reference.py``.  Only ``tests/`` may import this."""
import copy
from collections import deque
from typing import Dict, List

import numpy as np
import torch

ROLLOUT, BOUNDARY = 0, 1


def modify_adr_param(param, direction, prm, param_limit=None):
    """adr_vec_task.py:612-660."""
    op, delta = prm["delta_style"], prm["delta"]
    if direction == "up":
        new_val = param + delta if op == "additive" else param * delta
        if param_limit is not None:
            new_val = min(new_val, param_limit)
    else:
        new_val = param - delta if op == "additive" else param / delta
        if param_limit is not None:
            new_val = max(new_val, param_limit)
    return new_val, abs(new_val - param) > 1e-9


class AdrOracle:
    def __init__(self, adr_params: Dict, num_envs: int, queue_len: int, thr_low: float, thr_high: float):
        self.adr_params = copy.deepcopy(adr_params)
        for prm in self.adr_params.values():
            prm["range"] = list(prm["init_range"])
            prm.setdefault("limits", [None, None])
            prm.setdefault("delta_style", "additive")
        self.keys = list(self.adr_params.keys())
        self.queues: List[deque] = [deque(maxlen=queue_len) for _ in range(2 * len(self.keys))]
        self.queue_len, self.thr_low, self.thr_high = queue_len, thr_low, thr_high
        self.worker_types = torch.zeros(num_envs, dtype=torch.long)
        self.adr_modes = torch.zeros(num_envs, dtype=torch.long)
        self.npd = 0.0

    def update(self, rand_env_mask: torch.Tensor, objective: torch.Tensor) -> None:
        total_nats = 0.0
        for n, name in enumerate(self.keys):                      # order is immaterial without clear_other_queues
            prm = self.adr_params[name]
            low_idx, high_idx = 2 * n, 2 * n + 1
            w_low = (self.worker_types == BOUNDARY) & (self.adr_modes == low_idx)
            w_high = (self.worker_types == BOUNDARY) & (self.adr_modes == high_idx)
            self.queues[low_idx].extend(objective[rand_env_mask & w_low].numpy().tolist())
            self.queues[high_idx].extend(objective[rand_env_mask & w_high].numpy().tolist())
            lq, hq = self.queues[low_idx], self.queues[high_idx]
            mean_low = np.mean(lq) if len(lq) > 0 else 0.0
            mean_high = np.mean(hq) if len(hq) > 0 else 0.0
            lo, hi = prm["range"]
            limits, init = prm["limits"], prm["init_range"]
            nlo, nhi = prm.get("next_limits", [None, None])
            ch_lo = ch_hi = False
            if len(lq) >= self.queue_len:
                if mean_low < self.thr_low:
                    lo, ch_lo = modify_adr_param(lo, "up", prm, param_limit=init[0])
                elif mean_low > self.thr_high:
                    lo, ch_lo = modify_adr_param(lo, "down", prm, param_limit=limits[0])
                if ch_lo:
                    lq.clear()
                    self.worker_types[w_low] = ROLLOUT
            if len(hq) >= self.queue_len:
                if mean_high < self.thr_low:
                    hi, ch_hi = modify_adr_param(hi, "down", prm, param_limit=init[1])
                elif mean_high > self.thr_high:
                    hi, ch_hi = modify_adr_param(hi, "up", prm, param_limit=limits[1])
                if ch_hi:
                    hq.clear()
                    self.worker_types[w_high] = ROLLOUT
            if ch_lo or nlo is None:
                nlo, _ = modify_adr_param(lo, "down", prm, param_limit=limits[0])
            if ch_hi or nhi is None:
                nhi, _ = modify_adr_param(hi, "up", prm, param_limit=limits[1])
            prm["range"] = [lo, hi]
            if not prm["delta"] < 1e-9:
                total_nats += np.log(max(hi - lo, 1e-3))
            prm["next_limits"] = [nlo, nhi]
        self.npd = total_nats / len(self.adr_params)


def make_params(n_params: int, seed: int) -> Dict:
    """A synthetic ADR parameter table with both delta styles, limits on some parameters and one disabled (delta = 0)."""
    g = np.random.RandomState(seed)
    out = {}
    for i in range(n_params):
        mult = i % 3 == 1
        lo, hi = float(g.uniform(0.2, 1.0)), float(g.uniform(1.5, 3.0))
        prm = {"init_range": [lo, hi], "delta": 1.1 if mult else (0.0 if i == n_params - 1 else 0.05),
               "delta_style": "multiplicative" if mult else "additive"}
        if i % 2 == 0:
            prm["limits"] = [lo * 0.5, hi * 1.5]
        out[f"p{i}"] = prm
    return out
