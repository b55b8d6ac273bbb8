"""GPU: ``ha_observer_step`` / ``ha_observer_summary`` through ``handarm_b200.rlgames_adapter.RLGPUAlgoObserver`` against the
observer oracle (itself pinned to the reference's class): running sums and ring contents bit-exact (same fp32 adds, same
order of appends), mean / min / max within 1e-6 relative (float64 mean of fp32 values, summation order differs)."""
import math

import pytest
import torch

from oracle import observer_oracle as oo

pytestmark = pytest.mark.gpu


def _ring_as_deque(obs, k):
    """The device ring in deque order (oldest -> newest)."""
    cap, appended = obs.games_to_track, obs.episodes_finished()
    ring = obs.ring[k].cpu()
    if appended <= cap:
        return ring[:appended].tolist()
    head = appended % cap
    return torch.cat([ring[head:], ring[:head]]).tolist()


@pytest.mark.parametrize("n_envs,n_terms,cap,steps,p,use_mask", [
    (64, 3, 100, 30, 0.05, False), (1000, 8, 20, 25, 0.1, True), (33, 1, 5, 40, 0.5, False), (32768, 8, 100, 6, 0.02, True),
    (200000, 2, 100, 3, 0.01, True),        # more envs than one wave of 256-thread chunks: several sub-chunks per CTA
    (4096, 8, 100, 4, 0.0, True),           # nobody finishes
])
def test_observer_matches_oracle(n_envs, n_terms, cap, steps, p, use_mask):
    from handarm_b200.rlgames_adapter import RLGPUAlgoObserver
    keys, case = oo.make_case(n_envs, n_terms, steps, seed=11, done_prob=p)
    want = oo.ObserverOracle(cap)
    obs = RLGPUAlgoObserver(games_to_track=cap)
    for terms, dones in case:
        want.process_infos({"episode_cumulative": terms}, dones.nonzero(as_tuple=False))
        block = torch.stack([terms[k] for k in keys]).cuda()              # rows of one block, as the task exposes them
        infos = {"episode_cumulative": {k: block[i] for i, k in enumerate(keys)}, "scalars": {"tol": 0.1}}
        if use_mask:
            infos["dones"] = dones.cuda()
            obs.process_infos(infos, None)
        else:
            obs.process_infos(infos, dones.cuda().nonzero(as_tuple=False))
    torch.cuda.synchronize()
    n_done = sum(int(d.sum()) for _, d in case)
    assert obs.episodes_finished() == n_done
    for i, k in enumerate(keys):
        assert torch.equal(obs.episode_cumulative[k].cpu(), want.episode_cumulative[k]), k
        if n_done:
            assert _ring_as_deque(obs, i) == list(want.episode_cumulative_avg[k]), k
    got = obs.episode_cumulative_summary()
    if n_done:
        for k, (mean, mn, mx) in want.summary().items():
            assert math.isclose(got[k][0], mean, rel_tol=1e-6, abs_tol=1e-6)
            assert got[k][1] == pytest.approx(mn, rel=1e-7) and got[k][2] == pytest.approx(mx, rel=1e-7)
    else:
        assert all(math.isnan(v[0]) for v in got.values())
    assert "scalars/tol" in obs.direct_info


def test_observer_on_task_extras():
    """Driven by the task itself: extras['episode_cumulative'] rows are read in place, dones = reset_buf."""
    from handarm_b200.rlgames_adapter import RLGPUAlgoObserver
    from tests import kuka_harness as kh
    from oracle import kuka_state as ks
    n = 2048
    task = kh.make_task(n, device="cuda:0")
    kh.load_state(task, ks.make_state(n, seed=3, fresh=False))
    obs = RLGPUAlgoObserver(games_to_track=50)
    want = oo.ObserverOracle(50)
    n_done = 0
    for s in range(4):
        task.pre_physics_step(ks.make_actions(n, s).cuda())
        task.post_physics_step()
        infos = dict(task.extras)
        infos["dones"] = task.reset_buf
        n_done += int(task.reset_buf.sum())
        obs.process_infos(infos, None)
        want.process_infos({"episode_cumulative": {k: v.cpu().clone() for k, v in task.extras["episode_cumulative"].items()}},
                           task.reset_buf.cpu().nonzero(as_tuple=False))
    assert getattr(obs, "_stacked", None) is None, "the task's term block should be read in place"
    for k in want.episode_cumulative:
        assert torch.equal(obs.episode_cumulative[k].cpu(), want.episode_cumulative[k])
    assert obs.episodes_finished() == n_done


def test_observer_print_hooks():
    """after_print_stats: episode_cumulative scalars appear if and only if episodes finished since the last print (the reference's
    new_finished_episodes flag, rlgames_utils.py:196-202), the env's log_data is flushed (:206-214), and a games_to_track that
    after_init learns later re-sizes the device rings."""
    from types import SimpleNamespace
    from handarm_b200.rlgames_adapter import RLGPUAlgoObserver

    class Writer:
        def __init__(self):
            self.rows = []

        def add_scalar(self, tag, value, step):
            self.rows.append((tag, float(value), step))

    n, k = 256, 3
    env = SimpleNamespace(log_data={"custom/x": 1.5})
    algo = SimpleNamespace(writer=Writer(), games_to_track=7, vec_env=SimpleNamespace(env=env), device="cuda:0")
    obs = RLGPUAlgoObserver(games_to_track=100)
    terms = {f"t{i}": torch.full((n,), float(i + 1), device="cuda") for i in range(k)}
    dones = torch.zeros(n, dtype=torch.int64, device="cuda")
    obs.process_infos({"episode_cumulative": terms, "dones": dones}, None)       # sized with the constructor's window ...
    obs.after_init(algo)                                                         # ... then rl_games tells the real one
    assert obs.games_to_track == 7
    obs.process_infos({"episode_cumulative": terms, "dones": dones}, None)
    assert obs.ring.shape == (k, 7)
    obs.after_print_stats(10, 1, 0.5)
    tags = [t for t, _, _ in algo.writer.rows]
    assert not any(t.startswith("episode_cumulative") for t in tags), "nothing finished yet"
    assert "custom/x/frame" in tags and "custom/x/iter" in tags and "custom/x/time" in tags
    algo.writer.rows.clear()
    dones[5] = 1
    obs.process_infos({"episode_cumulative": terms, "dones": dones}, None)
    obs.after_print_stats(20, 2, 1.0)
    got = {t: v for t, v, _ in algo.writer.rows}
    assert got["episode_cumulative/t0"] == 2.0 and got["episode_cumulative_max/t2_max"] == 6.0      # two steps summed since the re-size
    assert obs.new_finished_episodes is False
    algo.writer.rows.clear()
    dones.zero_()
    obs.process_infos({"episode_cumulative": terms, "dones": dones}, None)
    obs.after_print_stats(30, 3, 1.5)
    assert not any(t.startswith("episode_cumulative") for t, _, _ in algo.writer.rows), "no new finished episodes since the last print"
