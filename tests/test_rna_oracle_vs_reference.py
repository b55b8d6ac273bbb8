"""CPU: the RNA restatement against the reference's own ``RandomNetworkAdversary`` module (``rna_util.py``) with the same
weights and masks: softmax output bit-exact."""
import importlib.util
import os

import pytest
import torch

from oracle import reference_loader as rl
from oracle import rna_oracle as ro

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference")]


def _ref_module():
    path = os.path.join(rl.REFERENCE_ROOT, "isaacgymenvs", "utils", "rna_util.py")
    spec = importlib.util.spec_from_file_location("ref_rna_util", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def test_restatement_matches_reference_module():
    n = 64
    ref = _ref_module().RandomNetworkAdversary(num_envs=n, in_dims=23, out_dims=16, softmax_bins=32, device="cpu")
    w = {k: v.detach().clone() for k, v in ref.state_dict().items()}
    w["dropout_masks1"], w["dropout_masks2"] = ref.dropout_masks1.clone(), ref.dropout_masks2.clone()
    x = torch.randn(n, 23, generator=torch.Generator().manual_seed(1))
    with torch.no_grad():
        want = ref(x)
    got = torch.softmax(ro.logits(w, x), dim=-1)
    assert torch.equal(want, got)
    # the seeded weight factory has the reference's shapes
    mine = ro.make_weights(n, 3)
    for k, v in w.items():
        assert mine[k].shape == v.shape, k
