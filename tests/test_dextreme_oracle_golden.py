"""The Dextreme restatement against the committed vectors generated from the reference's AllegroHandDextremeADR class
(oracle/make_golden_dextreme.py).  Runs anywhere (no reference tree needed)."""
import pytest
import torch

from oracle import dextreme_oracle as dx
from tests import golden_io as gio

PRE = ["reset_buf", "reset_goal_buf", "progress_buf", "hold_count_buf", "apply_reset_buf", "successes", "best_rotation_dist", "prev_targets",
       "cur_targets", "prev_prev_targets", "prev_actions", "prev_actions_queue", "rb_forces", "random_force_prob", "obs_object_pose",
       "obs_object_pose_freq", "actions", "dof_delta", "root_state", "dof_state", "goal_states"]
POST = PRE + ["rew_buf", "consecutive_successes", "object_pose", "object_pose_wrt_wrist", "timeout_buf", "episode_cumulative", "obs_rows",
              "obs_out", "frame"]


def test_fixtures_exist():
    assert len(gio.dextreme_fixture_names()) >= 3


@pytest.mark.parametrize("name", gio.dextreme_fixture_names())
def test_restatement_reproduces_reference_golden(name):
    c, st, steps = gio.load_dextreme_fixture(name)
    for i, s in enumerate(steps):
        lists = dx.pre_step(st, c, s["actions"], s["rnd"])
        for k in PRE:
            assert torch.equal(st[k].to(s["pre"][k].dtype), s["pre"][k]), f"{name} step {i} pre {k}"
        assert torch.equal(lists["root_rows"], s["pre"]["root_rows"])
        assert torch.equal(lists["hand_rows"], s["pre"]["hand_rows"])
        dx.post_step(st, c, s["prnd"])
        for k in POST:
            assert torch.equal(st[k].to(s["post"][k].dtype), s["post"][k]), f"{name} step {i} post {k}"
