"""CPU, world_size 2 over gloo: the N>1 path of the episode-statistics reduction (``handarm_b200/stats.py``).

Each rank owns a block of envs; the whole-job statistics must equal what one rank would have computed over the
concatenated blocks: SUM for sums / counts, MAX for [max, -min], means re-derived from the reduced sums."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from handarm_b200 import _lib, stats as hstats

C = _lib.CONSTS


def _rank_stats(rank: int, n: int) -> torch.Tensor:
    g = torch.Generator().manual_seed(100 + rank)
    rew, tobj = torch.randn(n, generator=g) * 10, torch.rand(n, generator=g) * 5
    prev = torch.randint(0, 9, (n,), generator=g).float()
    flags = (torch.rand(4, n, generator=g) < 0.1).float()
    s = torch.zeros(C["HA_STATS_LEN"])
    s[C["HA_KUKA_STAT_N"]] = n
    s[C["HA_KUKA_STAT_SUM_REWARD"]] = rew.sum()
    s[C["HA_KUKA_STAT_SUM_RESETS"]], s[C["HA_KUKA_STAT_SUM_SUCCESS_EVENTS"]] = flags[0].sum(), flags[1].sum()
    s[C["HA_KUKA_STAT_SUM_TIMEOUTS"]], s[C["HA_KUKA_STAT_SUM_LIFTED"]] = flags[2].sum(), flags[3].sum()
    s[C["HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES"]] = prev.sum()
    s[C["HA_KUKA_STAT_SUM_TRUE_OBJECTIVE"]] = tobj.sum()
    s[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]], s[C["HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE"]] = tobj.max(), -tobj.min()
    return hstats.derive(s)


def _worker(rank: int, world: int, port: int, sizes, out_dir: str):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        mine = _rank_stats(rank, sizes[rank])
        red = hstats.all_reduce_stats(mine)
        finish, handles = hstats.all_reduce_stats(mine, async_op=True)
        for h in handles:
            h.wait()
        red_async = finish()
        assert torch.equal(red, red_async)
        torch.save({"mine": mine, "reduced": red}, os.path.join(out_dir, f"r{rank}.pt"))
    finally:
        dist.destroy_process_group()


def test_two_rank_stats_all_reduce(tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    sizes = [8192, 4096]                                 # ragged shards
    mp.spawn(_worker, args=(2, port, sizes, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = torch.load(tmp_path / "r0.pt"), torch.load(tmp_path / "r1.pt")
    assert torch.equal(r0["reduced"], r1["reduced"])
    red, a, b = r0["reduced"], r0["mine"], r1["mine"]
    n = sum(sizes)
    assert red[C["HA_KUKA_STAT_N"]] == n
    for k in ("SUM_REWARD", "SUM_RESETS", "SUM_SUCCESS_EVENTS", "SUM_TIMEOUTS", "SUM_LIFTED", "SUM_PREV_EP_SUCCESSES", "SUM_TRUE_OBJECTIVE"):
        i = C["HA_KUKA_STAT_" + k]
        assert torch.isclose(red[i], a[i] + b[i], rtol=1e-6), k
    assert red[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]] == max(a[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]], b[C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"]])
    assert red[C["HA_KUKA_STAT_MIN_TRUE_OBJECTIVE"]] == min(a[C["HA_KUKA_STAT_MIN_TRUE_OBJECTIVE"]], b[C["HA_KUKA_STAT_MIN_TRUE_OBJECTIVE"]])
    assert torch.isclose(red[C["HA_KUKA_STAT_MEAN_REWARD"]], (a[C["HA_KUKA_STAT_SUM_REWARD"]] + b[C["HA_KUKA_STAT_SUM_REWARD"]]) / n, rtol=1e-6)
    # success rate: whole-job success events per env-step
    summ = hstats.summary(red)
    assert abs(summ["success_events_per_env"] - float(a[C["HA_KUKA_STAT_SUM_SUCCESS_EVENTS"]] + b[C["HA_KUKA_STAT_SUM_SUCCESS_EVENTS"]]) / n) < 1e-7


def test_single_process_is_identity():
    s = _rank_stats(0, 1000)
    assert torch.equal(hstats.all_reduce_stats(s), s)
