"""ha_stats_exchange: every rank stores its statistics into every peer's buffer and reduces what arrived during the
previous exchange.  Several ranks are emulated on one GPU (their exchange buffers are plain tensors of the same device and their
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
    for t in range(8):                                    # more rounds than ring slots
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


def test_missing_peer_sets_status_instead_of_hanging():
    dev = torch.device("cuda:0")
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=dev) for _ in range(2)]
    r0 = PeerStatsExchange(0, 2, bufs, dev)
    v = torch.ones(L, device=dev)
    r0.exchange(v)
    r0.exchange(v)            # rank 1 never pushed exchange 1: bounded wait, then status
    torch.cuda.synchronize()
    assert r0.status() == 1


def test_argument_checks():
    lib = _lib.load()
    assert lib.ha_stats_exchange_bytes(0) == 0 and lib.ha_stats_exchange_bytes(17) == 0
    assert lib.ha_stats_exchange_bytes(8) == 3 * 8 * L * 4 + 3 * 8 * 8
