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


class PeerStatsExchange:
    """Whole-box statistics without a collective launch: ``ha_stats_exchange`` stores this rank's partials into
    every peer's exchange buffer over NVLink and reduces the newest exchange that all ranks have published (wait-free).

    ``peer_buffers``: one ``[ha_stats_exchange_bytes(world) / 4]`` float32 tensor per rank, this rank's own included, all
    addressable from this device -- the views of a ``torch.distributed._symmetric_memory`` allocation (``connect``), or plain
    tensors of one device when several ranks are emulated on one GPU (tests).  ``exchange(stats)`` returns the whole-box partials
    of the newest complete exchange (normally the previous call's: one exchange late, like the side-stream all-reduce it
    replaces); until there is one it returns the rank's own statistics.
    """

    def __init__(self, rank: int, world: int, peer_buffers, device):
        import ctypes
        self._ctypes = ctypes
        self._lib = _lib.load()
        self.rank, self.world, self.device = rank, world, torch.device(device)
        assert len(peer_buffers) == world
        need = int(self._lib.ha_stats_exchange_bytes(world)) // 4
        for t in peer_buffers:
            assert t.dtype == torch.float32 and t.numel() >= need and t.is_contiguous()
        self._peer_buffers = list(peer_buffers)
        self._ptrs = torch.tensor([t.data_ptr() for t in peer_buffers], dtype=torch.int64, device=self.device)
        self._seq = torch.zeros(4, dtype=torch.int64, device=self.device)
        self._status = torch.zeros(1, dtype=torch.int32, device=self.device)
        self._reduced = torch.zeros(C["HA_STATS_LEN"], dtype=torch.float32, device=self.device)
        self._first = True
        x = _lib.HaStatsExchange()
        x.peer_bufs = self._ptrs.data_ptr()
        x.rank, x.world = rank, world
        x.max_mask = sum(1 << i for i in _MAX_SLOTS)
        x.interval = 1
        x.seq = self._seq.data_ptr()
        x.status = self._status.data_ptr()
        self._x = x

    @staticmethod
    def buffer_numel(world: int) -> int:
        return int(_lib.load().ha_stats_exchange_bytes(world)) // 4

    @classmethod
    def connect(cls, group=None, device=None) -> "PeerStatsExchange":
        """One exchange buffer per rank in symmetric memory (CUDA IPC handles exchanged over the process group's store)."""
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem
        group = group if group is not None else dist.group.WORLD
        rank, world = dist.get_rank(group), dist.get_world_size(group)
        device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        n = cls.buffer_numel(world)
        buf = symm_mem.empty(n, dtype=torch.float32, device=device)
        buf.zero_()
        hdl = symm_mem.rendezvous(buf, group)
        peers = [buf if r == rank else hdl.get_buffer(r, (n,), torch.float32) for r in range(world)]
        torch.cuda.synchronize(device)
        hdl.barrier()                      # every rank's buffer is zeroed before anyone stores into it
        self = cls(rank, world, peers, device)
        self._handle = hdl
        return self

    def exchange(self, stats: torch.Tensor) -> torch.Tensor:
        assert stats.dtype == torch.float32 and stats.numel() >= C["HA_STATS_LEN"] and stats.device == self.device
        ct = self._ctypes
        stream = ct.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
        if self._first:
            self._reduced.copy_(stats[:C["HA_STATS_LEN"]])
            self._first = False
        _lib.check(self._lib.ha_stats_exchange(ct.byref(self._x), ct.c_void_p(stats.data_ptr()), ct.c_void_p(self._reduced.data_ptr()), stream))
        return self._reduced

    def reduced(self) -> torch.Tensor:
        """Derived whole-box vector of the last completed exchange."""
        return derive(self._reduced)

    def attach(self, task, interval: int = 1) -> None:
        """Run the exchange INSIDE the task's post-physics kernel from now on (``HaKukaBuffers.xchg``): no launch of its own.  Every
        ``interval``-th ``post_physics_step`` publishes the statistics vector the preceding ``pre_physics_step`` folded and updates
        ``reduced()`` (``interval`` must be the same on every rank)."""
        ct = self._ctypes
        if getattr(self, "_x_dev", None) is None or self._x.interval != int(interval):
            self._x.interval = int(interval)
            raw = ct.string_at(ct.addressof(self._x), ct.sizeof(self._x))
            self._x_dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.device)
        task._buffers.xchg = self._x_dev.data_ptr()
        task._buffers.xchg_reduced = self._reduced.data_ptr()
        task._stats_exchange = self          # keeps the device struct alive

    @staticmethod
    def detach(task) -> None:
        task._buffers.xchg = None
        task._buffers.xchg_reduced = None

    def status(self) -> int:
        return int(self._status.item())
