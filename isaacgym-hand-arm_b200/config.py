"""Task-config loader: Hydra-style composition of the task nodes into the dict a task constructor reads.

The reference composes its configs with Hydra + OmegaConf (``isaacgymenvs/__init__.py:9-57``,
``cfg/config.yaml``).  Neither is installed here, so this module implements the small subset the
five hot-path task configs use.  The nodes come from one of two places:

* by default ``cfg/task_tables.json`` -- the parsed VALUES of the reference's config nodes (keys, ``defaults`` lists,
  interpolation strings; transcribed by ``tools/transcribe_cfg.py``, the YAML files themselves are not shipped);
* when ``HANDARM_CFG_DIR`` (or ``set_cfg_dir``) names a Hydra tree -- e.g. a user's own, possibly edited,
  ``isaacgymenvs/cfg`` -- the unmodified YAML files under it.

Supported Hydra / OmegaConf features:

* ``defaults:`` composition inside ``cfg/task`` (``- AllegroKuka`` pulls a sibling file,
  ``- env: reorientation`` merges ``cfg/task/env/reorientation.yaml`` under the ``env`` key,
  ``_self_`` fixes the merge order);
* ``${a.b}`` absolute and ``${..a}`` relative interpolations;
* the custom resolvers registered by the reference: ``eq``, ``contains``, ``if``,
  ``resolve_default`` (``isaacgymenvs/__init__.py:9-13``).

The result is the plain nested ``dict`` that ``omegaconf_to_dict(cfg.task)`` hands to the task
constructors (``rlgames_utils.py:113-121``).
"""
from __future__ import annotations

import copy
import json
import os
import re
from typing import Any, Dict, List, Optional

import yaml

TABLE_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "cfg", "task_tables.json")
_cfg_dir: Optional[str] = os.environ.get("HANDARM_CFG_DIR") or None
_table: Optional[dict] = None


def set_cfg_dir(path: Optional[str]) -> None:
    """Read the nodes from the Hydra YAML tree at ``path`` (``.../isaacgymenvs/cfg``); ``None`` returns to the built-in table."""
    global _cfg_dir
    if path is not None and not os.path.isfile(os.path.join(path, "config.yaml")):
        raise FileNotFoundError(f"{path!r} is not a Hydra config tree (no config.yaml)")
    _cfg_dir = path


def _merge(dst: dict, src: dict) -> dict:
    for k, v in src.items():
        if isinstance(v, dict) and isinstance(dst.get(k), dict):
            _merge(dst[k], v)
        else:
            dst[k] = copy.deepcopy(v)
    return dst


def _load_node(rel: str) -> dict:
    """One config node by its path relative to the tree root without extension (``task/AllegroKuka``, ``task/env/throw``)."""
    global _table
    if _cfg_dir is not None:
        path = os.path.join(_cfg_dir, rel + ".yaml")
        if not os.path.isfile(path):
            raise KeyError(f"no config node {rel!r} under {_cfg_dir}")
        with open(path) as f:
            return yaml.safe_load(f) or {}
    if _table is None:
        with open(TABLE_PATH) as f:
            _table = json.load(f)["nodes"]
    if rel not in _table:
        raise KeyError(f"no config node {rel!r} in {os.path.basename(TABLE_PATH)} (known: {sorted(_table)})")
    return copy.deepcopy(_table[rel])


def _compose_task(name: str, subdir: str = "task", overrides: Optional[Dict[str, str]] = None) -> dict:
    """Hydra-style composition of ``cfg/<subdir>/<name>.yaml`` honouring its ``defaults`` list."""
    overrides = overrides or {}
    raw = _load_node(subdir + "/" + name)
    defaults = raw.pop("defaults", None)
    if not defaults:
        return raw
    out: dict = {}
    saw_self = False
    for item in defaults:
        if item == "_self_":
            _merge(out, raw)
            saw_self = True
        elif isinstance(item, str):
            _merge(out, _compose_task(item, subdir, overrides))
        elif isinstance(item, dict):
            (group, choice), = item.items()
            choice = overrides.get(group, choice)
            sub = _compose_task(choice, subdir + "/" + group)
            _merge(out.setdefault(group, {}), sub)
        else:
            raise ValueError(f"unsupported defaults entry {item!r} in {name}.yaml")
    if not saw_self:  # Hydra >= 1.1 appends _self_ last when it is not listed
        _merge(out, raw)
    return out


# ---- interpolation -------------------------------------------------------------------------

_RESOLVERS = {
    "eq": lambda x, y: str(x).lower() == str(y).lower(),
    "contains": lambda x, y: str(x).lower() in str(y).lower(),
    "if": lambda pred, a, b: a if pred else b,
    "resolve_default": lambda default, arg: default if arg == "" else arg,
}


def _parse_scalar(tok: str) -> Any:
    tok = tok.strip()
    if len(tok) >= 2 and tok[0] == tok[-1] and tok[0] in "\"'":
        return tok[1:-1]
    try:
        return yaml.safe_load(tok) if tok != "" else ""
    except yaml.YAMLError:
        return tok


def _split_args(s: str) -> List[str]:
    args, depth, cur = [], 0, ""
    for ch in s:
        if ch == "," and depth == 0:
            args.append(cur)
            cur = ""
            continue
        if ch == "{":
            depth += 1
        elif ch == "}":
            depth -= 1
        cur += ch
    args.append(cur)
    return args


class _Resolver:
    def __init__(self, root: dict):
        self.root = root

    def lookup(self, ref: str, path: List[str]) -> Any:
        if ref.startswith("."):
            ups = len(ref) - len(ref.lstrip("."))
            # '.' = sibling of the current key, each extra dot climbs one level
            base = path[: len(path) - ups]
            keys = base + [k for k in ref.lstrip(".").split(".") if k]
        else:
            keys = ref.split(".")
        node: Any = self.root
        for i, k in enumerate(keys):
            node = node[k]
            node = self.resolve(node, keys[: i + 1])
        return node

    def _eval(self, expr: str, path: List[str]) -> Any:
        expr = expr.strip()
        m = re.match(r"^(\w+):(.*)$", expr, flags=re.S)
        if m and m.group(1) in _RESOLVERS:
            args = [self.resolve(_parse_scalar(a) if "${" not in a else a.strip(), path) for a in _split_args(m.group(2))]
            return _RESOLVERS[m.group(1)](*args)
        return self.lookup(expr, path)

    def resolve(self, value: Any, path: List[str]) -> Any:
        if not isinstance(value, str) or "${" not in value:
            return value
        # innermost-first substitution
        while True:
            m = None
            depth_start = []
            for i, ch in enumerate(value):
                if value.startswith("${", i):
                    depth_start.append(i)
                elif ch == "}" and depth_start:
                    start = depth_start.pop()
                    m = (start, i)
                    break
            if m is None:
                return value
            start, end = m
            res = self._eval(value[start + 2: end], path)
            if start == 0 and end == len(value) - 1:
                return self.resolve(res, path)
            value = value[:start] + (res if isinstance(res, str) else repr(res) if not isinstance(res, bool) else str(res)) + value[end + 1:]

    def walk(self, node: Any, path: List[str]) -> Any:
        if isinstance(node, dict):
            return {k: self.walk(v, path + [k]) for k, v in node.items()}
        if isinstance(node, list):
            return [self.walk(v, path + [str(i)]) for i, v in enumerate(node)]
        return self.resolve(node, path)


def load_task_config(task: str, num_envs: Optional[int] = None, sim_device: str = "cuda:0", subtask: Optional[str] = None,
                     overrides: Optional[Dict[str, Any]] = None) -> dict:
    """Return the resolved task config dict (what ``omegaconf_to_dict(cfg.task)`` gives the reference).

    ``overrides`` is a flat ``{"env.episodeLength": 300}`` style mapping applied before
    interpolation, like Hydra command-line overrides ``task.env.episodeLength=300``.
    """
    root = _load_node("config")
    root.pop("defaults", None)
    root.pop("hydra", None)
    root["num_envs"] = "" if num_envs is None else int(num_envs)
    root["sim_device"] = sim_device
    root["task"] = _compose_task(task, "task", {"env": subtask} if subtask else None)
    root["train"] = {"params": {"config": {"name": task}}}
    for key, val in (overrides or {}).items():
        node = root["task"]
        parts = key.split(".")
        for k in parts[:-1]:
            node = node.setdefault(k, {})
        node[parts[-1]] = val
    res = _Resolver(root)
    return res.walk(root["task"], ["task"])


def available_tasks() -> List[str]:
    if _cfg_dir is not None:
        return sorted(os.path.splitext(f)[0] for f in os.listdir(os.path.join(_cfg_dir, "task")) if f.endswith(".yaml"))
    _load_node("config")
    return sorted(k[len("task/"):] for k in _table if k.startswith("task/") and k.count("/") == 1)
