"""Observation / action noise of domain randomisation on the streaming sm_100a kernel ``ha_dr_noise_apply``.

Host mirror of the non-physical half of ``VecTask.apply_randomizations`` (``isaacgymenvs/tasks/base/vec_task.py:684-754``): the
schedule arithmetic stays in Python (a handful of scalars), the closure body -- ``op(tensor, corr * s + m + white * s' + m')`` over the
whole ``[N, W]`` tensor with fresh white noise every call -- is one kernel launch reading the tensor and the persistent correlated
noise once and writing the result once (the reference: ``randn_like`` + 5 elementwise launches and 4 temporaries per call).

Reference quirk kept: the closures read the loop variable ``op`` late, so both closures of one ``apply_randomizations`` call combine
with the operation of the LAST non-physical parameter configured (``vec_task.py:689``).
"""
from __future__ import annotations

import ctypes
from typing import Dict, Optional

import torch

from . import _lib

C = _lib.CONSTS


class NoiseClosure:
    """``noise_lambda`` of ``self.dr_randomizations[name]`` (callable; ``.params`` carries the reference's dict entries)."""

    def __init__(self, name: str, distribution: str, operation: int, a: float, b: float, a_corr: float, b_corr: float, seed: int):
        self.name, self.distribution = name, distribution
        p = _lib.HaNoiseParams()
        p.abi_version = C["HA_ABI_VERSION"]
        p.distribution = C["HA_NOISE_GAUSSIAN"] if distribution == "gaussian" else C["HA_NOISE_UNIFORM"]
        p.operation = operation
        if distribution == "gaussian":
            p.corr_scale, p.corr_shift, p.white_scale, p.white_shift = b_corr, a_corr, b, a          # var_corr, mu_corr, var, mu
        else:
            p.corr_scale, p.corr_shift, p.white_scale, p.white_shift = b_corr - a_corr, a_corr, b - a, a
        p.rng_seed, p.rng_step, p.use_fast_rng = seed, 0, 1
        self._p = p
        self.corr: Optional[torch.Tensor] = None
        self.ext_white: Optional[torch.Tensor] = None          # parity mode: raw draws of the next call
        self.out: Optional[torch.Tensor] = None

    def __call__(self, tensor: torch.Tensor, out: Optional[torch.Tensor] = None, clip: float = 0.0) -> torch.Tensor:
        if tensor.device.type != "cuda":
            raise _lib.HandArmLibraryError("DR noise runs only on a CUDA device (no CPU fallback)")
        x = tensor if (tensor.dtype == torch.float32 and tensor.is_contiguous()) else tensor.float().contiguous()
        if self.corr is None or self.corr.shape != x.shape:
            self.corr = torch.randn_like(x)                     # drawn once, on first use (vec_task.py:712-715)
        if out is None:
            if self.out is None or self.out.shape != x.shape:
                self.out = torch.empty_like(x)
            out = self.out
        p = self._p
        p.clip = clip
        if self.ext_white is not None:
            p.use_fast_rng, white = 0, ctypes.c_void_p(self.ext_white.data_ptr())
        else:
            p.use_fast_rng, white = 1, ctypes.c_void_p(0)
        stream = ctypes.c_void_p(torch.cuda.current_stream(x.device).cuda_stream)
        _lib.check(_lib.load().ha_dr_noise_apply(ctypes.byref(p), ctypes.c_void_p(x.data_ptr()), ctypes.c_void_p(self.corr.data_ptr()),
                                                 ctypes.c_void_p(out.data_ptr()), ctypes.c_size_t(x.numel()), white, stream))
        p.rng_step += 1
        return out


def build_noise_closures(dr_params: Dict, last_step: int, seed: int = 0) -> Dict[str, Dict]:
    """``self.dr_randomizations`` entries for "observations" / "actions" (vec_task.py:684-754)."""
    out: Dict[str, Dict] = {}
    last_op = None
    for i, name in enumerate(("observations", "actions")):
        if name not in dr_params:
            continue
        prm = dr_params[name]
        dist, op_type = prm["distribution"], prm["operation"]
        sched_type = prm["schedule"] if "schedule" in prm else None
        sched_step = prm["schedule_steps"] if "schedule" in prm else None
        last_op = C["HA_NOISE_ADDITIVE"] if op_type == "additive" else C["HA_NOISE_SCALING"]
        if sched_type == "linear":
            s = 1.0 / sched_step * min(last_step, sched_step)
        elif sched_type == "constant":
            s = 0 if last_step < sched_step else 1
        else:
            s = 1
        a, b = prm["range"]
        ac, bc = prm.get("range_correlated", [0., 0.])
        if dist not in ("gaussian", "uniform"):
            continue
        if op_type == "additive":
            a, b, ac, bc = a * s, b * s, ac * s, bc * s
        elif op_type == "scaling":
            keep = 1.0 * (1.0 - s)
            if dist == "gaussian":                       # the variance is scaled, the mean interpolated towards 1
                a, b, ac, bc = a * s + keep, b * s, ac * s + keep, bc * s
            else:
                a, b, ac, bc = a * s + keep, b * s + keep, ac * s + keep, bc * s + keep
        names = ("mu", "var", "mu_corr", "var_corr") if dist == "gaussian" else ("lo", "hi", "lo_corr", "hi_corr")
        out[name] = dict(zip(names, (a, b, ac, bc)))
        out[name]["_build"] = (name, dist, a, b, ac, bc, seed * 2 + i)
    for entry in out.values():
        name, dist, a, b, ac, bc, sd = entry.pop("_build")
        entry["noise_lambda"] = NoiseClosure(name, dist, last_op, a, b, ac, bc, sd)
    return out
