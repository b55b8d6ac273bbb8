"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the random network adversary (``isaacgymenvs/utils/rna_util.py:124-143``
forward, ``allegro_hand_dextreme.py:868-876`` argmax / permute / gather).  Pinned bit-exactly against the reference's own module
in ``tests/test_rna_oracle_vs_reference.py``.  Only ``tests/`` may import this."""
from typing import Dict, Tuple

import torch
import torch.nn.functional as F


def make_weights(num_envs: int, seed: int) -> Dict[str, torch.Tensor]:
    """Seeded weights / masks with the reference's shapes and init distributions (kaiming_uniform_, Bernoulli masks)."""
    g = torch.Generator().manual_seed(seed)
    dims = [("fc1", 23, 512), ("fc1_1", 512, 512), ("fc2", 512, 1024), ("fc2_1", 1024, 1024), ("fc3", 1024, 512)]
    w = {}
    for name, fi, fo in dims:
        bound = (6.0 / fi) ** 0.5
        w[name + ".weight"] = (torch.rand(fo, fi, generator=g) * 2 - 1) * bound
        w[name + ".bias"] = (torch.rand(fo, generator=g) * 2 - 1) / fi ** 0.5
    p = torch.rand(2, generator=g)
    w["dropout_masks1"] = torch.bernoulli(torch.full((num_envs, 512), float(p[0])), generator=g)
    w["dropout_masks2"] = torch.bernoulli(torch.full((num_envs, 1024), float(p[1])), generator=g)
    return w


def logits(w: Dict[str, torch.Tensor], x: torch.Tensor) -> torch.Tensor:
    """rna_util.py:126-136 up to the view."""
    x = F.linear(x, w["fc1.weight"], w["fc1.bias"])
    x = F.relu(x)
    x = F.linear(x, w["fc1_1.weight"], w["fc1_1.bias"])
    x = w["dropout_masks1"] * x
    x = F.linear(x, w["fc2.weight"], w["fc2.bias"])
    x = F.relu(x)
    x = F.linear(x, w["fc2_1.weight"], w["fc2_1.bias"])
    x = w["dropout_masks2"] * x
    x = F.linear(x, w["fc3.weight"], w["fc3.bias"])
    return x.view(-1, 16, 32)


def perturbation(w, dof_pos: torch.Tensor, object_pose_wrt_wrist: torch.Tensor, discretised_dofs: torch.Tensor) -> Tuple[torch.Tensor, ...]:
    """allegro_hand_dextreme.py:868-876 -> (perturbation [N,16], chosen bins [N,16], logits [N,16,32])."""
    lg = logits(w, torch.cat([dof_pos, object_pose_wrt_wrist], axis=-1))
    soft = F.softmax(lg, dim=-1)
    inds = torch.argmax(soft, axis=-1)
    inds_t = torch.permute(inds, (1, 0))
    pert = torch.permute(torch.gather(discretised_dofs, 1, inds_t), (1, 0))
    return pert, inds, lg
