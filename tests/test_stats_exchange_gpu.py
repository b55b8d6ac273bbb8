"""ha_stats_exchange: every rank stores its statistics into every peer's buffer and reduces the newest exchange
that all ranks have published, without ever waiting.  Several ranks are emulated on one GPU (their exchange buffers are plain tensors of the same device and their
launches share a stream), which exercises the slot ring, the tags and the SUM / MAX split; the real multi-GPU run is checked
against nccl inside bench.py."""
import pytest
import torch

from handarm_b200 import _lib
from handarm_b200.stats import PeerStatsExchange, _MAX_SLOTS, derive

pytestmark = pytest.mark.gpu
L = _lib.CONSTS["HA_STATS_LEN"]


def expected(vectors):
    """Rank-order fp32 reduction, the association the kernel uses."""
    acc = vectors[0].clone()
    for v in vectors[1:]:
        s = acc + v
        s[_MAX_SLOTS] = torch.maximum(acc[_MAX_SLOTS], v[_MAX_SLOTS])
        acc = s
    return acc


@pytest.mark.parametrize("world", [1, 2, 4, 8, 16])
def test_emulated_ranks_reduce_previous_exchange(world):
    dev = torch.device("cuda:0")
    n = PeerStatsExchange.buffer_numel(world)
    bufs = [torch.zeros(n, device=dev) for _ in range(world)]
    ranks = [PeerStatsExchange(r, world, bufs, dev) for r in range(world)]
    g = torch.Generator().manual_seed(world)
    prev = None
    for t in range(40):                                   # more rounds than ring slots
        cur = [(torch.randn(L, generator=g) * 100).to(dev) for _ in range(world)]
        outs = [ranks[r].exchange(cur[r]).clone() for r in range(world)]
        torch.cuda.synchronize()
        for r in range(world):
            if prev is None:
                assert torch.equal(outs[r], cur[r])       # nothing to reduce yet: the rank's own vector
            else:
                assert torch.equal(outs[r], expected(prev)), f"round {t} rank {r}"
            assert ranks[r].status() == 0
        prev = cur
    want = derive(expected(prev))
    ranks[0].exchange(prev[0])                            # one more exchange exposes the last round
    torch.cuda.synchronize()
    assert torch.equal(ranks[0].reduced(), want)


def test_absent_peer_never_blocks():
    dev = torch.device("cuda:0")
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    r0 = PeerStatsExchange(0, 2, bufs, dev)
    v = torch.arange(L, device=dev, dtype=torch.float32)
    for _ in range(40):       # rank 1 never publishes anything: no complete exchange, nothing to wait for
        out = r0.exchange(v)
    torch.cuda.synchronize()
    assert torch.equal(out, v) and r0.status() == 0


def test_lagging_rank_reduces_consistent_exchanges():
    """Rank 0 runs several exchanges ahead of rank 1; each side reduces the newest exchange BOTH have published, vector for vector."""
    dev = torch.device("cuda:0")
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    r = [PeerStatsExchange(k, 2, bufs, dev) for k in range(2)]
    g = torch.Generator().manual_seed(9)
    vec = [[(torch.randn(L, generator=g) * 10).to(dev) for _ in range(12)] for _ in range(2)]      # vec[rank][exchange number - 1]
    for k in range(5):
        r[0].exchange(vec[0][k])                                  # rank 0: exchanges 1..5, rank 1 silent
    assert torch.equal(r[0]._reduced, vec[0][0])                  # still its own first vector
    out = r[1].exchange(vec[1][0]).clone()                        # rank 1, exchange 1: its own head is still 0
    assert torch.equal(out, vec[1][0])
    out = r[1].exchange(vec[1][1]).clone()                        # heads (5, 1): exchange 1 is complete
    assert torch.equal(out, expected([vec[0][0], vec[1][0]]))
    out = r[0].exchange(vec[0][5]).clone()                        # heads (5, 2): exchange 2
    assert torch.equal(out, expected([vec[0][1], vec[1][1]]))
    for k in range(2, 6):
        r[1].exchange(vec[1][k])                                  # rank 1 catches up to exchange 6
    out = r[1].exchange(vec[1][6]).clone()                        # heads (6, 6)
    assert torch.equal(out, expected([vec[0][5], vec[1][5]]))
    assert r[0].status() == 0 and r[1].status() == 0


def test_rank_a_whole_ring_behind_is_skipped_flagged_and_rejoins():
    dev = torch.device("cuda:0")
    ring = _lib.CONSTS["HA_XCHG_RING"]
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    r = [PeerStatsExchange(k, 2, bufs, dev) for k in range(2)]
    v0, v1 = torch.full((L,), 1.0, device=dev), torch.full((L,), 2.0, device=dev)
    r[1].exchange(v1)                                             # rank 1 publishes exchange 1 and stops
    for _ in range(ring + 4):
        out = r[0].exchange(v0).clone()                           # rank 0 runs on to exchange ring + 4
    assert torch.equal(out, expected([v0, v1])) and r[0].status() == 0      # exchange 1 was reduced while it was there
    late = r[1].exchange(v1).clone()                              # rank 1 comes back: its slot of exchange 1 on rank 0 is long gone ...
    assert torch.equal(late, v1)                                  # ... nothing to reduce, nothing torn,
    assert r[1].status() == 1                                     # and it rejoins at rank 0's exchange number instead of counting on
    assert int(r[1]._seq[0]) == ring + 4
    w0, w1 = torch.full((L,), 3.0, device=dev), torch.full((L,), 5.0, device=dev)
    r[0].exchange(w0)                                             # both publish the next number ...
    r[1].exchange(w1)
    out0 = r[0].exchange(w0).clone()                              # ... and the one after exposes it on both sides
    out1 = r[1].exchange(w1).clone()
    torch.cuda.synchronize()
    assert torch.equal(out0, expected([w0, w1])) and torch.equal(out1, expected([w0, w1]))


def test_exchange_inside_the_post_kernel():
    """HaKukaBuffers.xchg: two emulated ranks = two AllegroKuka tasks on one GPU.  Every post_physics_step publishes the statistics
    the preceding pre_physics_step folded (those of the previous post step); the reduction both sides see is their rank-order sum."""
    from tests import kuka_harness as kh
    dev = torch.device("cuda:0")
    n = 4096 + 64                                                # the last CTA of the post grid owns a ragged tile
    nb = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(nb, device=dev) for _ in range(2)]
    tasks, xs = [], []
    for k in range(2):
        t = kh.make_task(n, device="cuda:0", snapshots=2, state_layout="soa")
        t.sim.fill_kuka(goal_states=None)
        t.seed_rng(50 + k)
        x = PeerStatsExchange(k, 2, bufs, dev)
        x.attach(t)
        tasks.append(t)
        xs.append(x)
    act = torch.zeros(n, 23, device=dev)
    folded = []                                                   # folded[s][k]: stats vector rank k's pre step s folded
    for s in range(6):
        row = []
        for k in range(2):
            tasks[k].pre_physics_step(act)
            row.append(tasks[k]._stats.clone())
            tasks[k].sim.simulate()
            tasks[k].post_physics_step()                          # publishes row[k] as exchange s + 1, reduces the newest complete one
        folded.append(row)
        torch.cuda.synchronize()
        if s >= 1:
            # a launch reduces before it publishes: both ranks see exchange s (the vectors of step s - 1)
            assert torch.equal(xs[0]._reduced, expected([folded[s - 1][0], folded[s - 1][1]])), s
            assert torch.equal(xs[1]._reduced, expected([folded[s - 1][0], folded[s - 1][1]])), s
        assert xs[0].status() == 0 and xs[1].status() == 0
    assert float(xs[1].reduced()[_lib.CONSTS["HA_KUKA_STAT_N"]]) == 2 * n
    PeerStatsExchange.detach(tasks[0])
    before = xs[0]._reduced.clone()
    tasks[0].pre_physics_step(act)
    tasks[0].post_physics_step()
    torch.cuda.synchronize()
    assert torch.equal(xs[0]._reduced, before)                    # detached: the kernel leaves the exchange alone


def test_exchange_inside_the_post_kernel_every_third_step():
    from tests import kuka_harness as kh
    dev = torch.device("cuda:0")
    n = 1024
    bufs = [torch.zeros(PeerStatsExchange.buffer_numel(1), device=dev)]
    t = kh.make_task(n, device="cuda:0", snapshots=2)
    t.sim.fill_kuka(goal_states=None)
    x = PeerStatsExchange(0, 1, bufs, dev)
    x.attach(t, interval=3)
    act = torch.zeros(n, 23, device=dev)
    for s in range(1, 10):
        t.step(act)
        torch.cuda.synchronize()
        assert int(x._seq[2]) == s and int(x._seq[0]) == s // 3      # post steps seen, exchanges published


def test_argument_checks():
    lib = _lib.load()
    assert lib.ha_stats_exchange_bytes(0) == 0 and lib.ha_stats_exchange_bytes(17) == 0
    ring = _lib.CONSTS["HA_XCHG_RING"]
    assert lib.ha_stats_exchange_bytes(8) == ring * 8 * L * 8 + 8 * 8      # tagged 8-byte words + head[world]
