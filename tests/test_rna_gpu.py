"""GPU: the random network adversary (``handarm_b200.rna`` on ``csrc/ha_rna.cu`` + library GEMMs) against the CPU oracle (pinned to
the reference module) with the same weights and dropout masks.

Tolerances: logits within 1e-4 relative to the row scale (fp32 GEMMs, summation order differs from the CPU's); the chosen bin is
an argmax, so it must agree wherever the oracle's top-2 margin exceeds that tolerance (checked), and the disagreements that
remain must pick a bin whose oracle logit is within the tolerance of the maximum."""
import pytest
import torch

from oracle import rna_oracle as ro

pytestmark = pytest.mark.gpu


def _load(net, w):
    for name in ("fc1", "fc1_1", "fc2", "fc2_1", "fc3"):
        getattr(net, name).weight.copy_(w[name + ".weight"])
        getattr(net, name).bias.copy_(w[name + ".bias"])
    net.dropout_masks1.copy_(w["dropout_masks1"])
    net.dropout_masks2.copy_(w["dropout_masks2"])
    net._wt = None


# (n, impl).  The persistent tcgen05 kernel runs grid = min(tiles, SMs) CTAs over 128 x 256 tiles: up to 4096 envs every CTA owns at
# most one tile; 5000 envs (40 row tiles: 80 / 160 / 80 tiles per layer, 160 % 148 != 0: twelve CTAs take a second tile) and
# cfg 5's 16384 envs (128 row tiles: 256 / 512 tiles, 2-4 per CTA) exercise the TMEM accumulator ping-pong, the operand ring running
# through tile boundaries and the mbarrier phase wrap.
CASES = [(n, impl) for impl in ("tc", "fp32") for n in (1, 33, 1000, 4096)] + [(5000, "tc"), (16384, "tc"), (16384 + 77, "tc")]


@pytest.mark.parametrize("n,impl", CASES)
def test_rna_matches_oracle(n, impl):
    from handarm_b200.rna import RandomNetworkAdversary
    g = torch.Generator().manual_seed(5)
    w = ro.make_weights(n, seed=9)
    dof_pos = torch.rand(n, 16, generator=g) * 1.5 - 0.3
    dof_vel = torch.randn(n, 16, generator=g)
    pose = torch.cat([torch.randn(n, 3, generator=g) * 0.1, torch.nn.functional.normalize(torch.randn(n, 4, generator=g), dim=-1)], dim=-1)
    lo, hi = -0.47 - torch.rand(16, generator=g) * 0.1, 1.6 + torch.rand(16, generator=g) * 0.1
    disc = torch.stack([torch.linspace(float(lo[i]), float(hi[i]), steps=32) for i in range(16)])
    want, want_idx, want_logits = ro.perturbation(w, dof_pos, pose, disc)

    net = RandomNetworkAdversary(n, 23, 16, 32, "cuda:0", impl=impl)
    _load(net, w)
    dof_state = torch.stack([dof_pos, dof_vel], dim=-1).contiguous().cuda()
    got = net.perturbation(dof_state, pose.cuda(), disc.cuda()).cpu()
    got_idx = net.bins.cpu().long()
    scale = want_logits.abs().amax(dim=-1, keepdim=True).clamp_min(1e-6)
    tol = 1e-4
    if impl == "fp32":                 # the tensor-core path keeps the logits in TMEM: only the selection leaves the kernel
        got_logits = (net._logits + net.fc3.bias).cpu().view(n, 16, 32)
        assert ((got_logits - want_logits).abs() / scale).max() < tol
    top2 = want_logits.topk(2, dim=-1).values
    margin = (top2[..., 0] - top2[..., 1]) / scale[..., 0]
    agree = got_idx == want_idx
    assert agree[margin > 2 * tol].all()
    chosen = want_logits.gather(-1, got_idx.unsqueeze(-1))[..., 0]
    assert ((top2[..., 0] - chosen) / scale[..., 0] <= 2 * tol).all()
    assert torch.equal(got[agree], want[agree])
    # reference surface: softmax probabilities
    probs = net(torch.cat([dof_pos, pose], dim=-1).cuda()).cpu()
    assert torch.allclose(probs, torch.softmax(want_logits, dim=-1), rtol=1e-3, atol=1e-6)


def test_rna_feeds_the_dextreme_pre_step():
    """The task runs the network itself when the config enables it (random_network_adversary.enable, the default) and hands the
    perturbation to the pre-physics kernel."""
    from oracle import dextreme_oracle as dx
    from tests.test_dextreme_parity_gpu import make_task
    task = make_task(512, dx.DexConsts())
    task.run_rna = True
    assert task.enable_rna and task.rna_network is not None
    task.step(torch.zeros(512, 16, device="cuda:0"))
    assert task.rna_perturbation is not None and task.rna_perturbation.shape == (512, 16)
    lo, hi = task.discretised_dofs[:, 0], task.discretised_dofs[:, -1]
    assert (task.rna_perturbation >= lo - 1e-6).all() and (task.rna_perturbation <= hi + 1e-6).all()


def test_rna_weights_are_resampled_at_the_configured_frequency():
    """allegro_hand_dextreme.py:868 + adr_vec_task.py:939: last_step follows the frame count, so the adversary's weights and dropout
    masks are re-drawn every weight_sample_freq steps (the attribute used to stay at -1 here)."""
    from oracle import dextreme_oracle as dx
    from tests.test_dextreme_parity_gpu import make_task
    task = make_task(256, dx.DexConsts())
    task.run_rna = True
    task.adr_resample_on_step = True
    task.random_adversary_weight_sample_freq = 3
    calls = []
    orig = task.rna_network._refresh
    task.rna_network._refresh = lambda: (calls.append(task.last_step), orig())[1]
    for _ in range(8):
        task.step(torch.zeros(256, 16, device="cuda:0"))
    assert task.last_step == 7
    assert calls == [3, 6]


def test_rna_tc_is_deterministic():
    """The persistent tcgen05 chain twice on the same input at cfg 5's size: identical bins and perturbations (a race in the mbarrier
    ring or between the TMEM accumulator buffers would show as run-to-run differences)."""
    from handarm_b200.rna import RandomNetworkAdversary
    n = 16384
    g = torch.Generator().manual_seed(11)
    w = ro.make_weights(n, seed=3)
    net = RandomNetworkAdversary(n, 23, 16, 32, "cuda:0", impl="tc")
    _load(net, w)
    dof_state = torch.randn(n, 16, 2, generator=g).cuda()
    pose = torch.randn(n, 7, generator=g).cuda()
    disc = torch.stack([torch.linspace(-0.5, 1.7, steps=32) for _ in range(16)]).cuda()
    outs = []
    for _ in range(3):
        outs.append((net.perturbation(dof_state, pose, disc).clone(), net.bins.clone()))
    torch.cuda.synchronize()
    for o in outs[1:]:
        assert torch.equal(o[0], outs[0][0]) and torch.equal(o[1], outs[0][1])
