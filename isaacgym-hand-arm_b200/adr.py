"""ADR range update (``ADRVecTask.adr_update``, ``isaacgymenvs/tasks/dextreme/adr_vec_task.py:760-918``) with the objective queues
on the device.

The reference extends ``2 * n_params`` host deques per call through boolean masks, gathers and one ``.cpu().numpy().tolist()``
sync per queue.  ``AdrRangeUpdater.update`` runs ``ha_adr_queue_step`` (one ordered compaction of the finished boundary workers, one
CTA per queue extending a device ring), reads ``[mean, length]`` of every queue in ONE device->host copy and keeps the scalar range
logic (thresholds, ``modify_adr_param``, next_limits, nats) on the host exactly as the reference has it; a boundary change clears
the queue (``appended[q] = 0``) and turns its workers into rollout workers with one masked device write.  The recycle / resample
tail of ``adr_update`` is ``ha_dex_adr_resample``.  ``clear_other_queues: True`` (not the shipped config) is not built."""
from __future__ import annotations

import ctypes
import math
from typing import Dict, Optional

import torch

from . import _lib

C = _lib.CONSTS


def modify_adr_param(param, direction, prm, param_limit=None):
    """adr_vec_task.py:612-660."""
    op, delta = prm["delta_style"], prm["delta"]
    if direction == "up":
        if op == "additive":
            new_val = param + delta
        else:
            assert delta > 1.0, "Must have delta>1 for multiplicative ADR update."
            new_val = param * delta
        if param_limit is not None:
            new_val = min(new_val, param_limit)
    else:
        if op == "additive":
            new_val = param - delta
        else:
            assert delta > 1.0, "Must have delta>1 for multiplicative ADR update."
            new_val = param / delta
        if param_limit is not None:
            new_val = max(new_val, param_limit)
    return new_val, abs(new_val - param) > 1e-9


class AdrRangeUpdater:
    def __init__(self, adr_params: Dict, num_envs: int, device, worker_types: torch.Tensor, adr_modes: torch.Tensor, queue_len: int,
                 thr_low: float, thr_high: float, clear_other_queues: bool = False):
        if torch.device(device).type != "cuda":
            raise _lib.HandArmLibraryError("AdrRangeUpdater needs a CUDA device: there is no CPU fallback")
        if clear_other_queues:
            raise NotImplementedError("adr.clear_other_queues: True is not built (the shipped DeXtreme config has it off)")
        self._lib = _lib.load()
        self.adr_params, self.keys = adr_params, list(adr_params.keys())          # shared with the task: ranges are updated in place
        for prm in adr_params.values():
            prm.setdefault("limits", [None, None])
            prm.setdefault("delta_style", "additive")
            assert prm["delta_style"] in ("additive", "multiplicative")
        self.num_envs, self.queue_len, self.thr_low, self.thr_high = num_envs, queue_len, thr_low, thr_high
        self.worker_types, self.adr_modes = worker_types, adr_modes
        Q = 2 * len(self.keys)
        self.n_queues = Q
        i64 = dict(dtype=torch.int64, device=device)
        self.queues = torch.zeros(Q, queue_len, dtype=torch.float32, device=device)
        self.appended = torch.zeros(Q, **i64)
        self._list = torch.zeros(num_envs, dtype=torch.int32, device=device)
        self._state = torch.zeros(4, **i64)
        self._block_counts = torch.zeros(C["HA_OBSERVER_MAX_CTAS"], **i64)
        self._summary = torch.zeros(Q, 2, dtype=torch.float64, device=device)
        self._summary_host = torch.zeros(Q, 2, dtype=torch.float64).pin_memory()
        b = _lib.HaAdrQueueBuffers()
        b.worker_types, b.adr_modes = worker_types.data_ptr(), adr_modes.data_ptr()
        b.queues, b.appended, b.list = self.queues.data_ptr(), self.appended.data_ptr(), self._list.data_ptr()
        b.state, b.block_counts, b.summary = self._state.data_ptr(), self._block_counts.data_ptr(), self._summary.data_ptr()
        self._b = b
        self.npd = 0.0
        self.last_summary = None

    def queue(self, q: int):
        """Queue ``q`` in deque order (oldest first) -- for tests and logging."""
        n = int(self.appended[q])
        ring = self.queues[q].cpu()
        if n <= self.queue_len:
            return ring[:n].tolist()
        h = n % self.queue_len
        return torch.cat([ring[h:], ring[:h]]).tolist()

    def update(self, reset_buf: torch.Tensor, apply_reset_buf: Optional[torch.Tensor], objective: torch.Tensor, all_envs: bool = False,
               extras: Optional[dict] = None, log: bool = False) -> bool:
        """One ``adr_update`` up to the recycle tail; returns True when a range or next_limit changed (the caller refreshes the
        table the resampling kernel reads).  ``reset_buf & ~apply_reset_buf`` (or every env) is ``rand_envs``."""
        b = self._b
        assert reset_buf.dtype == torch.int64 and objective.dtype == torch.float32 and objective.is_contiguous()
        b.reset_buf = reset_buf.data_ptr()
        b.apply_reset_buf = apply_reset_buf.data_ptr() if apply_reset_buf is not None else None
        b.objective = objective.data_ptr()
        stream = torch.cuda.current_stream(objective.device)
        _lib.check(self._lib.ha_adr_queue_step(self.num_envs, self.n_queues, self.queue_len, 1 if all_envs else 0, ctypes.byref(b),
                                               ctypes.c_void_p(stream.cuda_stream)))
        self._summary_host.copy_(self._summary, non_blocking=True)               # the one device->host copy of the update
        stream.synchronize()
        summ = self._summary_host.tolist()
        self.last_summary = summ
        total_nats, dirty = 0.0, False
        for n, name in enumerate(self.keys):                                      # (:775-779) the shuffle only matters with clear_other_queues
            prm = self.adr_params[name]
            low_idx, high_idx = 2 * n, 2 * n + 1
            (mean_low, len_low), (mean_high, len_high) = summ[low_idx], summ[high_idx]
            range_lower, range_upper = prm["range"]
            range_limits, init_range = prm["limits"], prm["init_range"]
            next_limit_lower, next_limit_upper = prm.get("next_limits", [None, None])
            changed_low = changed_high = False
            if len_low >= self.queue_len:                                         # (:819-840)
                if mean_low < self.thr_low:
                    range_lower, changed_low = modify_adr_param(range_lower, "up", prm, param_limit=init_range[0])
                elif mean_low > self.thr_high:
                    range_lower, changed_low = modify_adr_param(range_lower, "down", prm, param_limit=range_limits[0])
                if changed_low:
                    self._invalidate(low_idx)
            if len_high >= self.queue_len:                                        # (:842-861)
                if mean_high < self.thr_low:
                    range_upper, changed_high = modify_adr_param(range_upper, "down", prm, param_limit=init_range[1])
                elif mean_high > self.thr_high:
                    range_upper, changed_high = modify_adr_param(range_upper, "up", prm, param_limit=range_limits[1])
                if changed_high:
                    self._invalidate(high_idx)
            if changed_low or next_limit_lower is None:
                next_limit_lower, _ = modify_adr_param(range_lower, "down", prm, param_limit=range_limits[0])
                dirty = True
            if changed_high or next_limit_upper is None:
                next_limit_upper, _ = modify_adr_param(range_upper, "up", prm, param_limit=range_limits[1])
                dirty = True
            prm["range"] = [range_lower, range_upper]
            if not prm["delta"] < 1e-9:                                           # (:873-881)
                total_nats += math.log(max(range_upper - range_lower, 1e-3))
            prm["next_limits"] = [next_limit_lower, next_limit_upper]
            if extras is not None and (changed_low or changed_high or log):       # (:885-891)
                extras[f"adr/params/{name}/lower"] = range_lower
                extras[f"adr/params/{name}/upper"] = range_upper
                extras[f"adr/objective_perf/boundary/{name}/lower/value"] = mean_low
                extras[f"adr/objective_perf/boundary/{name}/lower/queue_len"] = int(len_low)
                extras[f"adr/objective_perf/boundary/{name}/upper/value"] = mean_high
                extras[f"adr/objective_perf/boundary/{name}/upper/queue_len"] = int(len_high)
        self.npd = total_nats / len(self.adr_params)
        if extras is not None and log:
            extras["adr/npd"] = self.npd
        return dirty

    def _invalidate(self, q: int) -> None:
        """The boundary moved: the queue restarts and its workers keep training as rollout workers (:836-840, :857-861)."""
        self.appended[q] = 0
        sel = (self.worker_types == C["HA_ADR_WORKER_BOUNDARY"]) & (self.adr_modes == q)
        self.worker_types.masked_fill_(sel, C["HA_ADR_WORKER_ROLLOUT"])
