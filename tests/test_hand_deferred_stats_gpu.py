"""Deferred statistics of the in-hand family: the post kernel leaves per-CTA slots, the fold runs at the head of the next pre step,
inside ``finalize_stats()`` or both.  Whatever the schedule, ``stats`` and the running ``consecutive_successes`` average
(shadow_hand.py:795-798: applied exactly once per post step) must come out bit-identical."""
import pytest
import torch

from oracle import hand_oracle as ho
from oracle import hand_state as hs
from tests import hand_harness as hh

pytestmark = pytest.mark.gpu


def run(n, consts, steps, schedule):
    task = hh.make_task(n, consts)
    hh.load_state(task, hs.make_state(n, 7, consts, fresh=False))
    g = torch.Generator(device="cpu").manual_seed(3)
    cons, stats = [], []
    for t in range(steps):
        task.ext_rand = hs.make_rnd(n, 20 + t, consts).cuda()
        task.pre_physics_step(hs.make_actions(n, 20 + t, consts).cuda())
        # short episodes with successes on the books: a share of the envs finishes every step, so the average moves every step
        task.progress_buf.copy_(torch.randint(0, consts.max_episode_length, (n,), generator=g).to(task.device))
        task.successes.copy_(torch.randint(0, 5, (n,), generator=g).float().to(task.device))
        task.post_physics_step()
        if schedule == "finalize":
            task.finalize_stats()
        elif schedule == "finalize_twice":
            task.finalize_stats()
            task._stats_pending = True      # a second fold of the same slots must not move the average
            task.finalize_stats()
        elif schedule == "mixed" and t % 3 == 0:
            task.finalize_stats()
        if schedule != "pre_only" or t == steps - 1:
            cons.append(task.consecutive_successes.cpu().clone())
            stats.append(task.stats.cpu().clone())
    torch.cuda.synchronize()
    return cons, stats


@pytest.mark.parametrize("n", [64, 1000, 4096])
def test_fold_schedules_agree(n):
    consts = ho.HandConsts(obs_type="full", max_episode_length=8, max_consecutive_successes=0)
    ref_cons, ref_stats = run(n, consts, 12, "finalize")
    assert float(ref_cons[-1]) != 0.0, "the case must move the average"
    assert len({float(c) for c in ref_cons}) > 3
    for schedule in ("finalize_twice", "mixed"):
        cons, stats = run(n, consts, 12, schedule)
        for a, b in zip(ref_cons, cons):
            assert torch.equal(a, b), schedule
        for a, b in zip(ref_stats, stats):
            assert torch.equal(a, b), schedule
    cons, stats = run(n, consts, 12, "pre_only")
    assert torch.equal(cons[-1], ref_cons[-1]) and torch.equal(stats[-1], ref_stats[-1])
