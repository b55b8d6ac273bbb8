"""ORACLE (test infrastructure, not product code) -- CPU restatement of the observation / action noise closures that
``VecTask.apply_randomizations`` builds (``vec_task.py:684-754``) and ``VecTask.step`` applies (``vec_task.py:398-400``, ``:427-429``).

Only ``tests/`` may import this.  Pinned against the reference's own closures by ``tests/test_dr_noise_oracle_vs_reference.py``.

Reference quirk kept on purpose: the closures read the loop variable ``op`` late (``vec_task.py:689`` / ``:720`` / ``:746``), so after one
``apply_randomizations`` call BOTH closures combine with the operation of the LAST non-physical parameter processed ("actions" when
both are configured).
"""
from __future__ import annotations

from typing import Dict, Optional

import torch

GAUSSIAN, UNIFORM = 0, 1
ADDITIVE, SCALING = 0, 1


def noise_specs(dr_params: Dict, last_step: int) -> Dict[str, Dict]:
    """The scalars of ``self.dr_randomizations[name]`` (schedule applied) plus the effective operation."""
    out: Dict[str, Dict] = {}
    last_op = None
    for name in ("observations", "actions"):
        if name not in dr_params:
            continue
        prm = dr_params[name]
        dist, op_type = prm["distribution"], prm["operation"]
        sched_type = prm["schedule"] if "schedule" in prm else None
        sched_step = prm["schedule_steps"] if "schedule" in prm else None
        last_op = ADDITIVE if op_type == "additive" else SCALING
        if sched_type == "linear":
            s = 1.0 / sched_step * min(last_step, sched_step)
        elif sched_type == "constant":
            s = 0 if last_step < sched_step else 1
        else:
            s = 1
        a, b = prm["range"]
        ac, bc = prm.get("range_correlated", [0., 0.])
        if dist == "gaussian":
            if op_type == "additive":
                a *= s; b *= s; ac *= s; bc *= s
            elif op_type == "scaling":
                b = b * s
                a = a * s + 1.0 * (1.0 - s)
                bc = bc * s
                ac = ac * s + 1.0 * (1.0 - s)
            out[name] = {"distribution": GAUSSIAN, "a": a, "b": b, "a_corr": ac, "b_corr": bc}
        elif dist == "uniform":
            if op_type == "additive":
                a *= s; b *= s; ac *= s; bc *= s
            elif op_type == "scaling":
                a = a * s + 1.0 * (1.0 - s)
                b = b * s + 1.0 * (1.0 - s)
                ac = ac * s + 1.0 * (1.0 - s)
                bc = bc * s + 1.0 * (1.0 - s)
            out[name] = {"distribution": UNIFORM, "a": a, "b": b, "a_corr": ac, "b_corr": bc}
    for spec in out.values():
        spec["operation"] = last_op          # late-bound ``op``
    return out


def apply_noise(x: torch.Tensor, spec: Dict, corr: torch.Tensor, white: torch.Tensor) -> torch.Tensor:
    """``noise_lambda(tensor)``: ``corr`` = the persistent N(0,1) tensor drawn on first use, ``white`` = this call's raw draw
    (N(0,1) for gaussian, U(0,1) for uniform)."""
    if spec["distribution"] == GAUSSIAN:
        c = corr * spec["b_corr"] + spec["a_corr"]                       # corr * var_corr + mu_corr
        noise = c + white * spec["b"] + spec["a"]                        # + randn * var + mu
    else:
        c = corr * (spec["b_corr"] - spec["a_corr"]) + spec["a_corr"]    # corr * (hi_corr - lo_corr) + lo_corr
        noise = c + white * (spec["b"] - spec["a"]) + spec["a"]          # + rand * (hi - lo) + lo
    return x + noise if spec["operation"] == ADDITIVE else x * noise
