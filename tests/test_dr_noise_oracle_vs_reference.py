"""The DR-noise restatement against the closures the reference's VecTask.apply_randomizations builds (vec_task.py:684-754)."""
import copy

import pytest
import torch

from oracle import dr_noise_oracle as dn
from oracle import reference_loader as rl

pytestmark = pytest.mark.skipif(not rl.reference_available(), reason="needs the reference tree")

CASES = {
    "gauss_add_linear": {"observations": {"range": [0.1, 0.02], "range_correlated": [0.05, 0.01], "operation": "additive", "distribution": "gaussian",
                                          "schedule": "linear", "schedule_steps": 10000}},
    "gauss_scale_const": {"observations": {"range": [1.0, 0.05], "range_correlated": [1.0, 0.02], "operation": "scaling", "distribution": "gaussian",
                                           "schedule": "constant", "schedule_steps": 3000}},
    "uniform_both": {"observations": {"range": [-0.02, 0.02], "range_correlated": [-0.01, 0.03], "operation": "additive", "distribution": "uniform"},
                     "actions": {"range": [0.9, 1.1], "operation": "scaling", "distribution": "uniform", "schedule": "linear", "schedule_steps": 20000}},
    "actions_gauss_add": {"actions": {"range": [0.0, 0.05], "range_correlated": [0.0, 0.015], "operation": "additive", "distribution": "gaussian"}},
}


def reference_closures(dr_params, last_step, n=16):
    mods = rl.load_reference()

    class Task(mods["vt"].VecTask):
        def create_sim(self):
            pass

        def pre_physics_step(self, actions):
            pass

        def post_physics_step(self):
            pass

    class Gym(rl.FakeGym):
        def get_frame_count(self, sim):
            return last_step

    env = object.__new__(Task)
    env.__dict__.update(gym=Gym(), sim=None, envs=[], first_randomization=True, last_step=-1, last_rand_step=-1, dr_randomizations={},
                        original_props={}, num_environments=n, randomize_buf=torch.zeros(n, dtype=torch.long),
                        reset_buf=torch.ones(n, dtype=torch.long), actor_params_generator=None, extern_actor_params={i: None for i in range(n)},
                        device="cpu")
    env.apply_randomizations(dict(copy.deepcopy(dr_params), frequency=10, actor_params={}))
    return env.dr_randomizations


@pytest.mark.parametrize("name", list(CASES))
@pytest.mark.parametrize("last_step", [0, 5000, 50000])
def test_noise_closures_are_bit_exact(name, last_step):
    dr = CASES[name]
    ref = reference_closures(dr, last_step)
    specs = dn.noise_specs(copy.deepcopy(dr), last_step)
    assert set(specs) == set(ref)
    gen = torch.Generator().manual_seed(3)
    for key, spec in specs.items():
        x = torch.randn(16, 23, generator=gen)
        corr, w1, w2 = torch.randn(16, 23, generator=gen), torch.randn(16, 23, generator=gen), torch.rand(16, 23, generator=gen)
        for white in ((w1, w2) if spec["distribution"] == dn.GAUSSIAN else (w2, w1.abs().clamp(max=1.0))):
            queue = [white] if "corr" in ref[key] else [corr, white]          # the first call draws the persistent corr tensor
            saved = (torch.randn_like, torch.rand_like)
            torch.randn_like = torch.rand_like = lambda t, **kw: queue.pop(0).clone()
            try:
                got_ref = ref[key]["noise_lambda"](x)
            finally:
                torch.randn_like, torch.rand_like = saved
            assert not queue
            assert torch.equal(dn.apply_noise(x, spec, corr, white), got_ref), f"{name}/{key} at step {last_step}"
