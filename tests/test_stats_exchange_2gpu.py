"""GPU x 2: the wait-free statistics exchange under TRUE concurrency -- two devices of one box storing into each other's buffers
over NVLink while both read (the one-GPU suite emulates the ranks serially on one stream; ADVICE r1 asked for this one).

Every rank publishes a vector that encodes (rank, exchange number) in every slot; whatever a rank reduces at any time must be the
exact reduction of ONE exchange number over both ranks -- a torn read (words of two exchanges mixed, a word without its tag) or a
lost update shows up as a value that belongs to no exchange.  Skipped on a box with one GPU (the driver's test box); run with
`gpurun --gpus 2 -- python -m pytest tests/test_stats_exchange_2gpu.py`.
"""
import pytest
import torch

from handarm_b200 import _lib
from handarm_b200.stats import PeerStatsExchange, _MAX_SLOTS

pytestmark = [pytest.mark.gpu,
              pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs with peer access")]
L = _lib.CONSTS["HA_STATS_LEN"]


def vector(rank, t):
    """slot i of rank r at exchange t: (r + 1) * 1000 * (i + 1) + t  (exact in fp32 for the sizes used here)"""
    i = torch.arange(L, dtype=torch.float32)
    return (rank + 1) * 1000.0 * (i + 1) + float(t)


def reduction(m):
    a, b = vector(0, m), vector(1, m)
    out = a + b
    out[_MAX_SLOTS] = torch.maximum(a[_MAX_SLOTS], b[_MAX_SLOTS])
    return out


@pytest.mark.parametrize("jitter", [0, 1])
def test_two_devices_exchange_concurrently(jitter):
    d0, d1 = torch.device("cuda:0"), torch.device("cuda:1")
    if not (torch.cuda.can_device_access_peer(0, 1) and torch.cuda.can_device_access_peer(1, 0)):
        pytest.skip("no peer access between the two devices")
    # torch enables peer access on the first peer copy in each direction
    torch.zeros(1, device=d0).copy_(torch.zeros(1, device=d1))
    torch.zeros(1, device=d1).copy_(torch.zeros(1, device=d0))
    n = PeerStatsExchange.buffer_numel(2)
    bufs = [torch.zeros(n, device=d0), torch.zeros(n, device=d1)]
    ranks = [PeerStatsExchange(0, 2, bufs, d0), PeerStatsExchange(1, 2, bufs, d1)]
    T = 3000
    stats = [torch.stack([vector(r, t) for t in range(1, T + 1)]).to(ranks[r].device) for r in range(2)]     # exchange numbers start at 1
    logs = [torch.zeros(T, L, device=ranks[r].device) for r in range(2)]
    streams = [torch.cuda.Stream(device=d0), torch.cuda.Stream(device=d1)]
    torch.cuda.synchronize(d0)
    torch.cuda.synchronize(d1)
    for t in range(T):
        for r in range(2):
            with torch.cuda.device(ranks[r].device), torch.cuda.stream(streams[r]):
                if jitter and (t * 7 + r * 3) % 11 == 0:
                    torch.cuda._sleep(20000 + 3000 * ((t + r) % 5))        # de-phase the two devices
                out = ranks[r].exchange(stats[r][t])
                logs[r][t].copy_(out, non_blocking=True)
    torch.cuda.synchronize(d0)
    torch.cuda.synchronize(d1)
    if (ranks[0].status() | ranks[1].status()) & 1:
        pytest.skip("a rank was lapped by a whole ring and rejoined (its exchange numbers jumped): the encoding below does not apply")
    want = torch.stack([reduction(m) for m in range(1, T + 1)])                                          # [T, L], row m-1 = exchange m
    for r in range(2):
        got = logs[r].cpu()
        own_first = vector(r, 1)
        last_m, reduced_any = 0, 0
        for t in range(T):
            row = got[t]
            if torch.equal(row, own_first):               # nothing complete yet: the rank's own first vector
                assert last_m == 0, f"rank {r} step {t}: fell back to its own vector after reducing exchange {last_m}"
                continue
            m = (float(row[0]) - 3000.0) / 2.0            # slot 0 is a SUM slot: 1000 + 2000 + 2 m
            assert m == int(m) and 1 <= m <= t + 1, f"rank {r} step {t}: slot 0 = {float(row[0])} belongs to no exchange"
            m = int(m)
            assert torch.equal(row, want[m - 1]), f"rank {r} step {t}: slots of different exchanges mixed (exchange {m})"
            assert m >= last_m, f"rank {r} step {t}: exchange number went back {last_m} -> {m}"
            last_m = m
            reduced_any += 1
        assert reduced_any > T // 2, f"rank {r}: only {reduced_any} of {T} calls saw a complete exchange"
        assert last_m >= T - 8, f"rank {r}: the last reduced exchange is {last_m} of {T}"
        assert ranks[r].status() & ~2 == 0
