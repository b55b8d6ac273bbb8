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


def test_rank_a_whole_ring_behind_is_skipped_and_flagged():
    dev = torch.device("cuda:0")
    ring = _lib.CONSTS["HA_XCHG_RING"]
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    r = [PeerStatsExchange(k, 2, bufs, dev) for k in range(2)]
    v0, v1 = torch.full((L,), 1.0, device=dev), torch.full((L,), 2.0, device=dev)
    r[1].exchange(v1)                                             # rank 1 publishes exchange 1 and stops
    for _ in range(ring + 4):
        out = r[0].exchange(v0).clone()
    assert r[0].status() == 0 and torch.equal(out, expected([v0, v1]))      # exchange 1 was reduced while it was there
    late = r[1].exchange(v1).clone()                              # exchange 2 of rank 1: rank 0's slot of exchange 1 is long gone
    assert torch.equal(late, v1) and r[1].status() == 0           # tags do not match: nothing reduced, nothing torn
    out = r[0].exchange(v0).clone()                               # newest common exchange is 2, a whole ring behind rank 0
    torch.cuda.synchronize()
    assert r[0].status() == 2 and torch.equal(out, expected([v0, v1]))


def test_argument_checks():
    lib = _lib.load()
    assert lib.ha_stats_exchange_bytes(0) == 0 and lib.ha_stats_exchange_bytes(17) == 0
    ring = _lib.CONSTS["HA_XCHG_RING"]
    assert lib.ha_stats_exchange_bytes(8) == ring * 8 * L * 4 + (ring + 1) * 8 * 8
