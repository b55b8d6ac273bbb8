"""CPU: the hand restatement against the golden vectors written by the reference's own ShadowHand /
AllegroHand classes (``oracle/make_golden_hand.py``).  Needs nothing but the committed fixtures."""
import pytest
import torch

from oracle import hand_oracle as ho
from tests import golden_io as gio


@pytest.mark.parametrize("name", gio.hand_fixture_names())
def test_hand_restatement_reproduces_reference_golden(name):
    c, st, steps = gio.load_hand_fixture(name)
    for i, s in enumerate(steps):
        lists = ho.pre_step(st, c, s["actions"], s["rnd"])
        for k, want in s["pre"].items():
            if k in ("root_rows", "hand_rows"):
                assert torch.equal(lists[k].to(torch.int32), want), f"{name} step {i} pre: {k}"
            else:
                assert torch.equal(st[k].reshape(-1), want.reshape(-1).to(st[k].dtype)), f"{name} step {i} pre: {k}"
        ho.post_step(st, c)
        for k, want in s["post"].items():
            assert torch.equal(st[k].reshape(-1), want.reshape(-1).to(st[k].dtype)), f"{name} step {i} post: {k}"


def test_fixtures_exist():
    assert len(gio.hand_fixture_names()) >= 5
