"""CPU, build container only: the restatement against the LIVE reference classes, bit for bit.

Skipped where ``/root/reference`` does not exist (the GPU box); there the committed golden vectors
(``tests/test_oracle_golden.py``) take over.
"""
import pytest
import torch

from oracle import kuka_oracle as ko, kuka_state as ks, reference_loader as rl

pytestmark = [pytest.mark.reference,
              pytest.mark.skipif(not rl.reference_available(), reason="reference tree not present on this machine")]


SUBTASKS = {
    "reorientation": ko.KukaConsts(),
    # env/regrasping.yaml, env/throw.yaml on top of AllegroKuka.yaml
    "regrasping": ko.KukaConsts(subtask="regrasping", success_steps=30, max_episode_length=300),
    "throw": ko.KukaConsts(subtask="throw", success_steps=5, max_episode_length=300, force_scale=0.0, success_tolerance=0.075,
                           initial_tolerance=0.075, target_tolerance=0.075),
}


@pytest.mark.parametrize("subtask", list(SUBTASKS))
@pytest.mark.parametrize("n,seed,fresh", [(257, 3, True), (257, 4, False), (1024, 5, False)])
def test_restatement_is_bit_exact_against_reference(n, seed, fresh, subtask):
    c = SUBTASKS[subtask]
    st = ks.make_state(n, seed=seed, c=c, fresh=fresh)
    st_ref = ks.clone_state(st)
    env = rl.make_reference_kuka(st_ref, c)
    for step in range(3):
        act, rnd = ks.make_actions(n, seed + step), ks.make_rnd(n, seed + step)
        lists = ko.pre_step(st, c, act, rnd)
        ko.post_step(st, c)
        env.gym.calls.clear()
        rl.reference_kuka_step(env, c, act, rnd, st_ref)
        out = rl.harvest(env, st_ref)
        for k, a in st.items():
            if k in out:
                b = out[k].reshape(a.shape)
                assert torch.equal(a.float(), b.float()), f"step {step} {k}: worst |d| {(a.float() - b.float()).abs().max():.3e}"
        rows = [args[2] for fn, args in env.gym.calls if fn == "set_actor_root_state_tensor_indexed"]
        assert torch.equal(torch.cat(rows) if rows else torch.zeros(0, dtype=torch.int32), lists["root_rows"])
        for k in ("reset_buf", "reset_goal_buf", "random_force_prob"):
            st_ref[k] = out[k]


def test_reference_quaternion_helpers_match_restatement():
    tju = rl.load_reference()["tju"]
    g = torch.Generator().manual_seed(0)
    q = torch.randn(512, 4, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    v = torch.randn(512, 3, generator=g)
    assert torch.equal(tju.quat_rotate(q, v), ko.quat_rotate(q, v))
    lo, hi = -torch.rand(7, generator=g) - 0.1, torch.rand(7, generator=g) + 0.1
    x = torch.randn(512, 7, generator=g)
    assert torch.equal(tju.unscale(x, lo, hi), ko.unscale(x, lo, hi))
    assert torch.equal(tju.scale(x, lo, hi), ko.scale(x, lo, hi))


@pytest.mark.parametrize("n_states", [40, 700, 3])
def test_initial_state_loading_matches_reference(n_states):
    """``loadInitialStates`` (allegro_kuka_base.py:1292-1313): consecutive recorded states for the reset envs in ascending order, the
    cursor wraps when the file runs out, nothing is loaded when one reset needs more states than the file holds."""
    c = ko.KukaConsts()
    n, seed = 257, 11
    st = ks.make_state(n, seed=seed, c=c, fresh=False)
    st_ref = ks.clone_state(st)
    g = torch.Generator().manual_seed(5)
    actors = st["root_state"].shape[0] // n
    root = torch.randn(n_states, actors, 13, generator=g)
    dof = torch.randn(n_states, c.n_dofs, 2, generator=g)
    initial = {"root": root, "dof": dof, "idx": 0}
    env = rl.make_reference_kuka(st_ref, c)
    env.should_load_initial_states = True
    env.initial_root_state_tensors, env.initial_dof_state_tensors = root.clone(), dof.clone()
    env.num_initial_states, env.initial_state_idx = n_states, 0
    for step in range(4):
        if step == 2:                                   # one step in which every env resets: more resets than a small file holds
            st["reset_buf"][:] = 1
            st_ref["reset_buf"][:] = 1
            env.reset_buf = st_ref["reset_buf"]
        act, rnd = ks.make_actions(n, seed + step), ks.make_rnd(n, seed + step)
        ko.pre_step(st, c, act, rnd, initial=initial)
        ko.post_step(st, c)
        rl.reference_kuka_step(env, c, act, rnd, st_ref)
        out = rl.harvest(env, st_ref)
        for k in ("dof_state", "root_state", "prev_targets", "cur_targets", "obs_buf", "reset_buf"):
            assert torch.equal(st[k].float().reshape(-1), out[k].float().reshape(-1)), f"step {step} {k}"
        assert initial["idx"] == env.initial_state_idx, f"step {step}: cursor"
        for k in ("reset_buf", "reset_goal_buf", "random_force_prob"):
            st_ref[k] = out[k]
