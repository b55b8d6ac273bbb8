"""Cross-rank reduction of the packed episode-statistics vector.

Environments shard embarrassingly over the GPUs of a box: one process per GPU, contiguous env
blocks, no data-path collective.  The only cross-env quantities are the few scalars the reference
logs per rank (``allegro_kuka_base.py:903-910``: mean of ``prev_episode_successes``, mean / min / max of
``true_objective``) plus success rate and mean reward.  ``ha_kuka_post_step`` leaves them in
``stats[0:HA_KUKA_N_PARTIAL_STATS]`` as per-rank partials: sums and counts, and ``[max, -min]`` so that one MAX
reduction covers both extremes.  ``all_reduce_stats`` combines them with one SUM and one MAX
all-reduce (NCCL over NVLink on the GPU box, gloo in the CPU tests) and re-derives the means.

The reference itself never reduces these across ranks; parity is defined per rank and the reduced
vector is an additional reported statistic.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

from . import _lib

C = _lib.CONSTS
_N_PART = C["HA_KUKA_N_PARTIAL_STATS"]
_MAX_SLOTS = [C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"], C["HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE"]]


def derive(stats: torch.Tensor) -> torch.Tensor:
    """Fill the derived slots (means, min) of a stats vector from its partial slots (same rule as the kernel's epilogue)."""
    out = stats.clone()
    n = out[C["HA_KUKA_STAT_N"]].clamp_min(1.0)
    out[C["HA_KUKA_STAT_MEAN_REWARD"]] = out[C["HA_KUKA_STAT_SUM_REWARD"]] / n
    out[C["HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES"]] = out[C["HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES"]] / n
    out[C["HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE"]] = out[C["HA_KUKA_STAT_SUM_TRUE_OBJECTIVE"]] / n
    out[C["HA_KUKA_STAT_MIN_TRUE_OBJECTIVE"]] = -out[C["HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE"]]
    return out


def all_reduce_stats(stats: torch.Tensor, group=None, async_op: bool = False):
    """Whole-job statistics from this rank's ``stats`` vector ([HA_STATS_LEN] float32, on the rank's device).

    Returns the reduced vector (a new tensor; ``stats`` is not modified).  With ``async_op=True`` returns
    ``(tensor, [work handles])``: the tensor is valid after the handles completed and ``finish`` ran.
    """
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        out = derive(stats)
        return (out, []) if async_op else out
    sums = stats[:_N_PART].clone()
    maxs = stats[_MAX_SLOTS].clone()
    sums[_MAX_SLOTS] = 0.0                                   # the MAX slots travel in their own reduction
    h1 = dist.all_reduce(sums, op=dist.ReduceOp.SUM, group=group, async_op=async_op)
    h2 = dist.all_reduce(maxs, op=dist.ReduceOp.MAX, group=group, async_op=async_op)

    def finish() -> torch.Tensor:
        out = stats.clone()
        out[:_N_PART] = sums
        out[_MAX_SLOTS] = maxs
        return derive(out)

    if async_op:
        return finish, [h1, h2]
    return finish()


def summary(stats: torch.Tensor) -> Dict[str, float]:
    """Host-side dict of the headline numbers (one device->host copy)."""
    s = stats.detach().float().cpu()
    n = max(float(s[C["HA_KUKA_STAT_N"]]), 1.0)
    return {
        "envs": n,
        "mean_reward": float(s[C["HA_KUKA_STAT_SUM_REWARD"]]) / n,
        "success_events_per_env": float(s[C["HA_KUKA_STAT_SUM_SUCCESS_EVENTS"]]) / n,
        "reset_rate": float(s[C["HA_KUKA_STAT_SUM_RESETS"]]) / n,
        "timeout_rate": float(s[C["HA_KUKA_STAT_SUM_TIMEOUTS"]]) / n,
        "lifted_rate": float(s[C["HA_KUKA_STAT_SUM_LIFTED"]]) / n,
        "mean_prev_episode_successes": float(s[C["HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES"]]) / n,
        "true_objective_mean": float(s[C["HA_KUKA_STAT_SUM_TRUE_OBJECTIVE"]]) / n,
        "true_objective_max": float(s[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]]),
        "true_objective_min": -float(s[C["HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE"]]),
    }
