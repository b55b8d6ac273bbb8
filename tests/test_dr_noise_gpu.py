"""GPU: the DR observation / action noise kernel (ha_dr_noise_apply) against the oracle of the reference's closures, and through
VecTask.step.  fp32 elementwise: 1e-5 relative, 2e-6 absolute."""
import copy

import pytest
import torch

from handarm_b200.dr_noise import build_noise_closures
from oracle import dr_noise_oracle as dn
from tests.test_dr_noise_oracle_vs_reference import CASES

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("shape", [(32768, 117), (33, 23), (1, 3)])
def test_noise_kernel_matches_oracle(name, shape):
    dr, last_step = CASES[name], 5000
    specs = dn.noise_specs(copy.deepcopy(dr), last_step)
    closures = build_noise_closures(copy.deepcopy(dr), last_step)
    assert set(specs) == set(closures)
    gen = torch.Generator().manual_seed(11)
    for key, spec in specs.items():
        x, corr = torch.randn(*shape, generator=gen), torch.randn(*shape, generator=gen)
        white = torch.randn(*shape, generator=gen) if spec["distribution"] == dn.GAUSSIAN else torch.rand(*shape, generator=gen)
        want = dn.apply_noise(x, spec, corr, white)
        fn = closures[key]["noise_lambda"]
        fn.corr, fn.ext_white = corr.cuda(), white.cuda()
        got = fn(x.cuda()).cpu()
        assert torch.allclose(got, want, rtol=1e-5, atol=2e-6), f"{name}/{key}: worst |d| = {(got - want).abs().max():.3e}"
        for k in ("mu", "var", "mu_corr", "var_corr", "lo", "hi", "lo_corr", "hi_corr"):
            if k in closures[key]:
                assert closures[key][k] == pytest.approx({"mu": spec["a"], "var": spec["b"], "mu_corr": spec["a_corr"], "var_corr": spec["b_corr"],
                                                          "lo": spec["a"], "hi": spec["b"], "lo_corr": spec["a_corr"], "hi_corr": spec["b_corr"]}[k])


def test_fast_mode_statistics_and_clip():
    dr = {"observations": {"range": [0.5, 0.1], "range_correlated": [0.0, 0.0], "operation": "additive", "distribution": "gaussian"}}
    fn = build_noise_closures(dr, 0, seed=3)["observations"]["noise_lambda"]
    x = torch.zeros(4096, 117, device="cuda:0")
    a, b = fn(x).clone(), fn(x).clone()
    assert abs(float(a.mean()) - 0.5) < 2e-3 and abs(float(a.std()) - 0.1) < 2e-3
    assert not torch.equal(a, b)                                   # fresh white noise every call
    assert float(fn(x, clip=0.55).max()) <= 0.55 + 1e-6


def test_step_applies_noise_closures():
    """VecTask.step applies the action closure before pre_physics_step and the observation closure after post_physics_step."""
    from tests import kuka_harness as kh
    task = kh.make_task(256, device="cuda:0")
    task.apply_randomizations({"frequency": 1, "observations": {"range": [0.0, 0.01], "operation": "additive", "distribution": "gaussian"},
                               "actions": {"range": [0.0, 0.02], "operation": "additive", "distribution": "gaussian"}})
    assert set(task.dr_randomizations) == {"observations", "actions"}
    obs, rew, done, info = task.step(torch.zeros(256, 23, device="cuda:0"))
    assert torch.isfinite(obs["obs"]).all() and obs["obs"].shape == (256, 117)
    assert float(task.actions.abs().max()) > 0.0                   # the zero action arrived with noise


def test_noised_observations_track_the_kernel_over_steps():
    """Several steps with observation noise on: the returned observations follow the kernel's fresh observations every step (the
    closure is applied in place on the tensor the kernels are bound to), stay within clip_obs, and the order is the reference's
    (clamp_obs -> noise -> clip, vec_task.py:427-432)."""
    from tests import kuka_harness as kh
    n = 512
    task = kh.make_task(n, device="cuda:0", snapshots=4)
    twin = kh.make_task(n, device="cuda:0", snapshots=4)          # same seeds, no noise
    for t in (task, twin):
        t.sim.fill_kuka(goal_states=None)
        t.seed_rng(5)
    task.apply_randomizations({"frequency": 1, "observations": {"range": [0.0, 0.02], "operation": "additive", "distribution": "gaussian"}})
    bound = task.obs_buf.data_ptr()
    act = torch.zeros(n, 23, device="cuda:0")
    prev = None
    for step in range(4):
        obs = task.step(act)[0]["obs"]
        clean = twin.step(act)[0]["obs"]
        assert task.obs_buf.data_ptr() == bound                   # never rebound
        assert float(obs.abs().max()) <= task.clip_obs + 1e-6
        inside = clean.abs() < task.clip_obs - 0.2                # away from the clip the difference is the noise alone
        diff = (obs - clean)[inside]
        assert abs(float(diff.mean())) < 2e-3 and abs(float(diff.std()) - 0.02) < 2e-3, (step, float(diff.mean()), float(diff.std()))
        if prev is not None:
            assert not torch.equal(prev, clean)                   # the simulator snapshots differ from step to step
        prev = clean.clone()


def test_step_can_return_a_fresh_observation_tensor():
    """obs_dict["obs"] aliases obs_buf by default (documented); return_obs_copy restores the reference's fresh tensor per call."""
    from tests import kuka_harness as kh
    task = kh.make_task(128, device="cuda:0", snapshots=2)
    task.sim.fill_kuka(goal_states=None)
    act = torch.zeros(128, 23, device="cuda:0")
    a = task.step(act)[0]["obs"]
    assert a.data_ptr() == task.obs_buf.data_ptr()
    task.return_obs_copy = True
    b = task.step(act)[0]["obs"]
    kept = b.clone()
    c = task.step(act)[0]["obs"]
    assert b.data_ptr() != task.obs_buf.data_ptr() and c.data_ptr() != b.data_ptr()
    assert torch.equal(b, kept)                                   # last step's observation did not mutate
