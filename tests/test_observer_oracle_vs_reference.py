"""CPU: the observer restatement (``oracle/observer_oracle.py``) against the reference's own ``RLGPUAlgoObserver``
(``rlgames_utils.py:129-205``) imported from ``/root/reference`` -- bit-exact deques and running sums."""
import types

import pytest
import torch

from oracle import observer_oracle as oo
from oracle import reference_loader as rl

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference")]


@pytest.mark.parametrize("n_envs,n_terms,cap,steps,p", [(64, 3, 100, 30, 0.05), (500, 8, 20, 25, 0.1), (33, 1, 5, 40, 0.5)])
def test_restatement_matches_reference_class(n_envs, n_terms, cap, steps, p):
    rlu = rl.load_rlgames_utils()
    ref = rlu.RLGPUAlgoObserver()
    ref.algo = types.SimpleNamespace(games_to_track=cap)
    mine = oo.ObserverOracle(cap)
    keys, case = oo.make_case(n_envs, n_terms, steps, seed=7, done_prob=p)
    for terms, dones in case:
        idx = dones.nonzero(as_tuple=False)
        ref.process_infos({"episode_cumulative": {k: v.clone() for k, v in terms.items()}}, idx)
        mine.process_infos({"episode_cumulative": terms}, idx)
    for k in keys:
        assert torch.equal(ref.episode_cumulative[k], mine.episode_cumulative[k])
        assert list(ref.episode_cumulative_avg[k]) == list(mine.episode_cumulative_avg[k])
    assert ref.new_finished_episodes == mine.new_finished_episodes
