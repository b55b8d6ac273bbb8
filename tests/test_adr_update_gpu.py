"""GPU: ``handarm_b200.adr.AdrRangeUpdater`` (``ha_adr_queue_step`` + the host range logic) against the ADR oracle, which is pinned
to the reference's ``adr_update``: ranges and next_limits equal as Python floats, queues equal and in deque order, worker types
bit-exact, over phases of good and bad performance that move boundaries both ways."""
import copy

import pytest
import torch

from oracle import adr_oracle as ao

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,n_params,qlen", [(600, 5, 16), (20000, 12, 8), (4096, 40, 256)])
def test_adr_update_matches_oracle(n, n_params, qlen):
    from handarm_b200.adr import AdrRangeUpdater
    params = ao.make_params(n_params, seed=2)
    want = ao.AdrOracle(params, n, qlen, 5.0, 20.0)
    mine_params = copy.deepcopy(params)
    for prm in mine_params.values():
        prm["range"] = list(prm["init_range"])
    wt = torch.zeros(n, dtype=torch.int64, device="cuda:0")
    md = torch.zeros(n, dtype=torch.int64, device="cuda:0")
    upd = AdrRangeUpdater(mine_params, n, "cuda:0", wt, md, qlen, 5.0, 20.0)
    g = torch.Generator().manual_seed(4)
    any_dirty = False
    for step in range(40):
        redo = torch.rand(n, generator=g) < 0.3
        new_wt = (torch.rand(n, generator=g) < 0.6).long()
        new_md = torch.randint(0, 2 * n_params, (n,), generator=g)
        want.worker_types[redo], want.adr_modes[redo] = new_wt[redo], new_md[redo]
        wt[redo.cuda()], md[redo.cuda()] = new_wt[redo].cuda(), new_md[redo].cuda()
        done = torch.rand(n, generator=g) < 0.25
        applied = torch.rand(n, generator=g) < 0.1                       # apply_reset_buf: those envs are not re-randomised
        hi_phase = (step // 8) % 3 != 2
        objective = torch.randint(21, 40, (n,), generator=g).float() if hi_phase else torch.randint(0, 4, (n,), generator=g).float()
        want.update(done & ~applied, objective)
        extras = {}
        any_dirty |= upd.update(done.long().cuda(), applied.long().cuda(), objective.cuda(), extras=extras, log=step % 10 == 0)
        for k in params:
            assert mine_params[k]["range"] == want.adr_params[k]["range"], (step, k)
            assert mine_params[k]["next_limits"] == want.adr_params[k]["next_limits"], (step, k)
        assert torch.equal(wt.cpu(), want.worker_types), step
        if step % 7 == 0 or step == 39:
            for q in range(2 * n_params):
                assert upd.queue(q) == list(want.queues[q]), (step, q)
        assert upd.npd == pytest.approx(want.npd, rel=1e-12)
    assert any_dirty
    if qlen <= 16:
        assert sum(mine_params[k]["range"] != list(params[k]["init_range"]) for k in params) >= 2


def test_adr_update_runs_inside_the_task_step():
    """The DeXtreme task runs the range update ahead of its resampling kernel when the config enables it (the default)."""
    from oracle import dextreme_oracle as dx
    from tests.test_dextreme_parity_gpu import make_task
    n = 2048
    task = make_task(n, dx.DexConsts())
    assert task.use_adr and task.update_adr_ranges
    task.adr_resample_on_step = True
    task.first_randomization = True
    for _ in range(4):
        task.step(torch.zeros(n, 16, device="cuda:0"))
    upd = task._adr_updater
    assert upd is not None and len(upd.last_summary) == 2 * task.num_adr_params
    lens = [l for _, l in upd.last_summary]
    assert all(0 <= l <= task.adr_queue_threshold_length for l in lens)
    for k, prm in task.adr_params.items():
        assert prm["next_limits"][0] is not None and prm["next_limits"][1] is not None
