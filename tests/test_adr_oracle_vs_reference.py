"""CPU: the ADR range-update restatement (``oracle/adr_oracle.py``) against the reference's own ``ADRVecTask.adr_update``
(``adr_vec_task.py:760-918``) driven on a bare instance -- ranges, next_limits, queues and worker types identical."""
import copy
from collections import deque

import pytest
import torch

from oracle import adr_oracle as ao
from oracle import reference_loader as rl

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference")]


def _bare_reference(params, n, qlen, lo, hi):
    adrv = rl.load_reference()["adrv"]
    cls = type("BareADR", (adrv.ADRVecTask,), {"pre_physics_step": lambda self, a: None, "post_physics_step": lambda self: None})
    t = object.__new__(cls)
    t.num_environments, t.device = n, "cpu"          # num_envs is a read-only property over num_environments
    t.update_adr_ranges, t.adr_clear_other_queues = True, False
    t.adr_params = copy.deepcopy(params)
    for prm in t.adr_params.values():
        prm["range"] = list(prm["init_range"])
        prm.setdefault("limits", [None, None])
        prm.setdefault("delta_style", "additive")
    t.num_adr_params = len(params)
    t.adr_queue_threshold_length, t.adr_objective_threshold_low, t.adr_objective_threshold_high = qlen, lo, hi
    t.adr_objective_queues = [deque(maxlen=qlen) for _ in range(2 * len(params))]
    t.worker_types = torch.zeros(n, dtype=torch.long)
    t.adr_modes = torch.zeros(n, dtype=torch.long)
    t.adr_tensor_values = {}
    t.extras, t.last_step = {}, 1
    t.adr_rollout_perf_last, t.adr_rollout_perf_alpha = None, 0.99
    t.recycle_envs = lambda env_ids: None            # the recycle / resample tail has its own kernel and tests
    return t


@pytest.mark.parametrize("n_params,qlen", [(5, 16), (12, 8)])
def test_adr_update_restatement_matches_reference(n_params, qlen):
    n = 600
    params = ao.make_params(n_params, seed=2)
    ref = _bare_reference(params, n, qlen, 5.0, 20.0)
    mine = ao.AdrOracle(params, n, qlen, 5.0, 20.0)
    g = torch.Generator().manual_seed(4)
    for step in range(60):
        # worker assignment as recycle_envs would leave it (random; the same on both sides)
        redo = torch.rand(n, generator=g) < 0.3
        wt = (torch.rand(n, generator=g) < 0.6).long()
        md = torch.randint(0, 2 * n_params, (n,), generator=g)
        for t in (ref, mine):
            t.worker_types[redo] = wt[redo]
            t.adr_modes[redo] = md[redo]
        done = torch.rand(n, generator=g) < 0.25
        hi_phase = (step // 8) % 3 != 2                     # phases of good (> threshold_high) and bad (< threshold_low) performance
        objective = torch.randint(21, 40, (n,), generator=g).float() if hi_phase else torch.randint(0, 4, (n,), generator=g).float()
        ref.adr_update(done.nonzero(as_tuple=False).squeeze(-1), objective)
        mine.update(done, objective)
        for k in params:
            assert ref.adr_params[k]["range"] == mine.adr_params[k]["range"], (step, k)
            assert ref.adr_params[k]["next_limits"] == mine.adr_params[k]["next_limits"], (step, k)
        assert torch.equal(ref.worker_types, mine.worker_types)
        for q in range(2 * n_params):
            assert list(ref.adr_objective_queues[q]) == list(mine.queues[q]), (step, q)
    changed = sum(ref.adr_params[k]["range"] != list(params[k]["init_range"]) for k in params)
    assert changed >= 2, "the case must exercise range changes"
