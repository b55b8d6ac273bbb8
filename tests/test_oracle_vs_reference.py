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
