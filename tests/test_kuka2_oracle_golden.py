"""CPU: the two-arm regrasping restatement against the golden vectors written by the reference's own
AllegroKukaTwoArmsRegrasping class (``oracle/make_golden_kuka2.py``)."""
import pytest
import torch

from oracle import kuka2_oracle as k2
from tests import golden_io as gio


@pytest.mark.parametrize("name", gio.kuka2_fixture_names())
def test_kuka2_restatement_reproduces_reference_golden(name):
    c, st, steps = gio.load_kuka2_fixture(name)
    for i, s in enumerate(steps):
        lists = k2.pre_step(st, c, s["actions"], s["rnd"])
        for k, want in s["pre"].items():
            got = lists[k].to(torch.int32) if k in ("root_rows", "hand_rows") else st[k]
            assert torch.equal(got.reshape(-1), want.reshape(-1).to(got.dtype)), f"{name} step {i} pre: {k}"
        info = k2.post_step(st, c)
        for k, want in s["post"].items():
            assert torch.equal(st[k].reshape(-1), want.reshape(-1).to(st[k].dtype)), f"{name} step {i} post: {k}"
        assert torch.equal(info["true_objective_mean"], s["extras"]["true_objective_mean"])


def test_kuka2_fixtures_exist():
    assert len(gio.kuka2_fixture_names()) >= 3
