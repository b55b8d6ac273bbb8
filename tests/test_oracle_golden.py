"""CPU: the oracle restatement (``oracle/kuka_oracle.py``) against the reference-generated golden vectors.

The fixtures were produced by running the reference's real ``AllegroKukaReorientation`` class
(``oracle/make_golden.py``).  Integer / flag / index outputs must be bit-exact.  Floating point
outputs are bit-exact on the machine that generated them; across CPUs torch's vectorised
sin/cos/log/exp may differ in the last ulp, so floats are compared at 1e-6 relative (1e-7 absolute).
"""
import pytest
import torch

from oracle import kuka_oracle as ko
from tests import golden_io as gio

EXACT = ["reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "timeout_buf", "near_goal_steps", "lifted_object"]
RTOL, ATOL = 1e-6, 1e-7


def compare(got, want, ctx, skip=()):
    for k, w in want.items():
        if k in skip or k in ("root_rows", "hand_rows"):
            continue
        g = got[k]
        if k in EXACT:
            assert torch.equal(g.long().reshape(-1), w.long().reshape(-1)), f"{ctx}{k}"
        else:
            g, w = g.float().reshape(-1), w.float().reshape(-1)
            assert torch.allclose(g, w, rtol=RTOL, atol=ATOL), f"{ctx}{k}: worst |d| = {(g - w).abs().max():.3e}"


def test_fixtures_present():
    assert len(gio.fixture_names()) >= 4


@pytest.mark.parametrize("name", gio.fixture_names())
def test_oracle_reproduces_reference_golden(name):
    c, st, steps = gio.load_fixture(name)
    for i, s in enumerate(steps):
        lists = ko.pre_step(st, c, s["actions"], s["rnd"])
        # episode_cumulative / obs / rew are written by the post step only
        compare(st, s["pre"], f"[{name} step {i} pre] ", skip=("obs_buf", "rew_buf", "episode_cumulative", "timeout_buf"))
        assert torch.equal(lists["root_rows"], s["pre"]["root_rows"]), "actor rows handed to set_actor_root_state_tensor_indexed"
        assert torch.equal(lists["hand_rows"], s["pre"]["hand_rows"]), "actor rows handed to set_dof_state_tensor_indexed"
        info = ko.post_step(st, c)
        compare(st, s["post"], f"[{name} step {i} post] ")
        ex = s["extras"]
        assert torch.allclose(info["successes_mean"], ex["successes"], rtol=1e-6)
        assert torch.allclose(info["true_objective_mean"], ex["true_objective_mean"], rtol=1e-6)
        assert torch.equal(info["true_objective_min"], ex["true_objective_min"])
        assert torch.equal(info["true_objective_max"], ex["true_objective_max"])


def test_interp_known_answers():
    """The reference's only in-tree known answers (allegro_kuka_utils.py:137-142): interp_0_1 at the
    initial, target and mid tolerance is 0, 1 and 0.5 -- restated by ``true_objective_coeffs``' offset."""
    initial, target = 0.075, 0.01
    assert abs(ko.true_objective_coeffs(initial, initial, target)[1] - 0.0) < 1e-5
    assert abs(ko.true_objective_coeffs(target, initial, target)[1] - 1.0) < 1e-5
    mid = (initial + target) / 2
    assert abs(ko.true_objective_coeffs(mid, initial, target)[1] - 0.5) < 1e-5
    assert ko.true_objective_coeffs(initial, initial, target)[0] == 0.01 and ko.true_objective_coeffs(target, initial, target)[0] == 1.0
