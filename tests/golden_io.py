"""Loads the reference-generated fixtures of ``tests/golden`` (made by ``oracle/make_golden.py``)."""
from __future__ import annotations

import glob
import os
from typing import Dict, List, Tuple

import numpy as np
import torch

from oracle import kuka_oracle as ko

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
INT_CONSTS = {"success_steps", "max_episode_length", "max_consecutive_successes"}


def fixture_names() -> List[str]:
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "kuka_*.npz")))


def load_fixture(name: str):
    """-> (consts, input state dict, [per-step dict(actions, rnd, pre={...}, post={...}, extras={...})])"""
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    over = {}
    for k in z.files:
        if k.startswith("consts/"):
            key = k.split("/", 1)[1]
            over[key] = str(z[k]) if z[k].dtype.kind in "US" else (int(z[k]) if key in INT_CONSTS else float(z[k]))
    c = ko.KukaConsts(**over)
    st = {k.split("/", 1)[1]: torch.from_numpy(z[k].copy()) for k in z.files if k.startswith("in/")}
    steps = []
    for i in range(int(z["meta/steps"])):
        d: Dict[str, object] = {"actions": torch.from_numpy(z[f"s{i}/actions"].copy()), "rnd": torch.from_numpy(z[f"s{i}/rnd"].copy())}
        for phase in ("pre", "post", "extras"):
            pre = f"s{i}/{phase}/"
            d[phase] = {k[len(pre):]: torch.from_numpy(np.asarray(z[k]).copy()) for k in z.files if k.startswith(pre)}
        steps.append(d)
    return c, st, steps


def hand_fixture_names() -> List[str]:
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "hand_*.npz")))


def load_hand_fixture(name: str):
    """-> (HandConsts, input state dict, [per-step dict(actions, rnd, pre={...}, post={...})])"""
    import json
    from oracle import hand_oracle as ho
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    d = json.loads(str(z["meta/consts_json"]))
    for k in ("actuated", "tip_bodies", "force_prob_range", "goal_displacement"):
        d[k] = tuple(d[k])
    c = ho.HandConsts(**d)
    st = {k.split("/", 1)[1]: torch.from_numpy(z[k].copy()) for k in z.files if k.startswith("in/")}
    steps = []
    for i in range(int(z["meta/steps"])):
        s: Dict[str, object] = {"actions": torch.from_numpy(z[f"s{i}/actions"].copy()), "rnd": torch.from_numpy(z[f"s{i}/rnd"].copy())}
        for phase in ("pre", "post"):
            pre = f"s{i}/{phase}/"
            s[phase] = {k[len(pre):]: torch.from_numpy(np.asarray(z[k]).copy()) for k in z.files if k.startswith(pre)}
        steps.append(s)
    return c, st, steps


def kuka2_fixture_names() -> List[str]:
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "kuka2_*.npz")))


def load_kuka2_fixture(name: str):
    """-> (Kuka2Consts, input state dict, [per-step dict(actions, rnd, pre={...}, post={...}, extras={...})])"""
    import json
    from oracle import kuka2_oracle as k2
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    d = json.loads(str(z["meta/consts_json"]))
    for k, v in list(d.items()):
        if isinstance(v, list):
            d[k] = tuple(tuple(x) if isinstance(x, list) else x for x in v)
    c = k2.Kuka2Consts(**d)
    st = {k.split("/", 1)[1]: torch.from_numpy(z[k].copy()) for k in z.files if k.startswith("in/")}
    steps = []
    for i in range(int(z["meta/steps"])):
        s: Dict[str, object] = {"actions": torch.from_numpy(z[f"s{i}/actions"].copy()), "rnd": torch.from_numpy(z[f"s{i}/rnd"].copy())}
        for phase in ("pre", "post", "extras"):
            pre = f"s{i}/{phase}/"
            s[phase] = {k[len(pre):]: torch.from_numpy(np.asarray(z[k]).copy()) for k in z.files if k.startswith(pre)}
        steps.append(s)
    return c, st, steps


def dextreme_fixture_names() -> List[str]:
    return sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "dextreme_*.npz")))


def load_dextreme_fixture(name: str):
    """-> (DexConsts, input state dict, [per-step dict(actions, rnd, prnd, pre={...}, post={...})])"""
    import json
    from oracle import dextreme_oracle as dx
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    d = json.loads(str(z["meta/consts_json"]))
    for k, v in list(d.items()):
        if isinstance(v, list):
            d[k] = tuple(v)
    c = dx.DexConsts(**d)
    st = {k.split("/", 1)[1]: torch.from_numpy(np.asarray(z[k]).copy()) for k in z.files if k.startswith("in/")}
    steps = []
    for i in range(int(z["meta/steps"])):
        s: Dict[str, object] = {k: torch.from_numpy(z[f"s{i}/{k}"].copy()) for k in ("actions", "rnd", "prnd")}
        for phase in ("pre", "post"):
            pre = f"s{i}/{phase}/"
            s[phase] = {k[len(pre):]: torch.from_numpy(np.asarray(z[k]).copy()) for k in z.files if k.startswith(pre)}
        steps.append(s)
    return c, st, steps
