"""CPU, build container only: the hand restatement (``oracle/hand_oracle.py``) against the reference's real
``ShadowHand`` / ``AllegroHand`` classes on identical inputs and identical random tables -- bit-exact."""
import pytest
import torch

from oracle import hand_oracle as ho
from oracle import hand_state as hs
from oracle import reference_loader as rl

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="/root/reference not present")]

CASES = {
    "shadow_full_state": ho.HandConsts(),
    "shadow_full": ho.HandConsts(obs_type="full"),
    "shadow_no_vel": ho.HandConsts(obs_type="full_no_vel"),
    "shadow_openai_asym_forces": ho.HandConsts(obs_type="openai", asymmetric=True, force_scale=1.0, max_consecutive_successes=50,
                                               fall_penalty=-50.0, act_moving_average=0.3),
    "shadow_relative": ho.HandConsts(obs_type="full", use_relative_control=True, reset_dof_vel_noise=0.1),
    "shadow_pen": ho.HandConsts(ignore_z_rot=True),        # objectType "pen": randomize_rotation_pen + doubled success tolerance
    "allegro_full_state": ho.allegro_consts(),
    "allegro_full": ho.allegro_consts(obs_type="full", force_scale=2.0, max_consecutive_successes=10, fall_penalty=-10.0),
    "allegro_no_vel_asym": ho.allegro_consts(obs_type="full_no_vel", asymmetric=True, act_moving_average=0.5),
}
KEYS = ["obs_buf", "states_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "successes",
        "consecutive_successes", "timeout_buf", "prev_targets", "cur_targets", "actions", "rb_forces", "random_force_prob", "goal_states",
        "root_state", "dof_state"]


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("fresh", [False, True])
def test_hand_restatement_is_bit_exact(name, fresh):
    c = CASES[name]
    n = 257
    st_ref = hs.make_state(n, 5, c, fresh=fresh)
    st = {k: v.clone() for k, v in st_ref.items()}
    env = rl.make_reference_hand(st_ref, c)
    for step in range(3):
        act, rnd = hs.make_actions(n, step, c), hs.make_rnd(n, step, c)
        rl.reference_hand_step(env, c, act, rnd, st_ref)
        st_ref = rl.harvest_hand(env, st_ref, c)
        lists = ho.pre_step(st, c, act, rnd)
        ho.post_step(st, c)
        rows = [a[2] for fn, a in env.gym.calls if fn == "set_actor_root_state_tensor_indexed"]
        env.gym.calls.clear()
        if rows:
            assert torch.equal(torch.sort(rows[-1].to(torch.int32)).values, lists["root_rows"])
        for k in KEYS:
            assert torch.equal(st[k].reshape(-1), st_ref[k].reshape(-1).to(st[k].dtype)), f"{name} step {step}: {k} differs"
