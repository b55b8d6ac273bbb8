"""CPU, build container only: the two-arm restatement (``oracle/kuka2_oracle.py``) against the reference's real
``AllegroKukaTwoArmsRegrasping`` / ``AllegroKukaTwoArmsReorientation`` classes on identical inputs and identical random tables --
bit-exact."""
import pytest
import torch

from oracle import kuka2_oracle as k2
from oracle import reference_loader as rl

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="/root/reference not present")]

KEYS = ["obs_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "successes", "prev_episode_successes",
        "true_objective", "prev_episode_true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
        "closest_fingertip_dist", "timeout_buf", "prev_targets", "cur_targets", "actions", "rb_forces", "random_force_prob", "goal_states",
        "root_state", "dof_state", "rewards_episode", "episode_cumulative"]


@pytest.mark.parametrize("fresh", [False, True])
@pytest.mark.parametrize("subtask", ["regrasping", "reorientation"])
@pytest.mark.parametrize("over", [{}, dict(success_steps=1, force_scale=0.0, act_moving_average=0.4, success_tolerance=0.04)])
def test_kuka2_restatement_is_bit_exact(fresh, over, subtask):
    c = k2.Kuka2Consts(subtask=subtask, **over)
    n = 193
    st_ref = k2.make_state(n, 7, c, fresh=fresh)
    st = {k: v.clone() for k, v in st_ref.items()}
    env = rl.make_reference_kuka2(st_ref, c)
    for step in range(3):
        act, rnd = k2.make_actions(n, step, c), k2.make_rnd(n, step, c)
        env.gym.calls.clear()
        rl.reference_kuka2_step(env, c, act, rnd, st_ref)
        st_ref = rl.harvest_kuka2(env, st_ref)
        lists = k2.pre_step(st, c, act, rnd)
        info = k2.post_step(st, c)
        rows = [a[2] for fn, a in env.gym.calls if fn == "set_actor_root_state_tensor_indexed"]
        if rows:
            assert torch.equal(rows[-1].to(torch.int32), lists["root_rows"])
        hands = [a[2] for fn, a in env.gym.calls if fn == "set_dof_state_tensor_indexed"]
        if hands:
            assert torch.equal(hands[-1].to(torch.int32), lists["hand_rows"])
        for k in KEYS:
            assert torch.equal(st[k].reshape(-1), st_ref[k].reshape(-1).to(st[k].dtype)), f"step {step}: {k} differs"
        assert torch.equal(info["true_objective_mean"], env.extras["true_objective_mean"])
