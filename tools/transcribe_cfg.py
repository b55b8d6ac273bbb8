#!/usr/bin/env python
"""Transcribe the VALUES of a Hydra task-config tree into ``handarm_b200/cfg/task_tables.json``.

The task constructors read their scalars from the dict Hydra composes out of ``isaacgymenvs/cfg`` (``config.yaml``,
``task/*.yaml``, ``task/env/*.yaml``).  The repo does not ship those files; it ships one JSON table with the parsed
content of each node (keys, values, ``defaults`` lists and ``${...}`` interpolation strings, in declaration order, no
comments), which ``handarm_b200.config`` composes and resolves exactly as it would the YAML tree.  A user who has the
reference checkout points ``HANDARM_CFG_DIR`` at its ``isaacgymenvs/cfg`` instead and the same loader reads the YAML.

Run here (where ``/root/reference`` exists) whenever the upstream tree changes:

    python tools/transcribe_cfg.py [--src /root/reference/isaacgymenvs/cfg]

``tests/test_config.py::test_table_matches_reference_tree`` re-checks the table against the tree when it is present.
"""
import argparse
import json
import os

import yaml

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(HERE, "..", "isaacgym-hand-arm_b200", "cfg", "task_tables.json")

# the hand / hand-arm task nodes of SURVEY.md 8a (+ the variants that inherit from them) and the env groups they select
TASKS = ["ShadowHand", "ShadowHandTest", "ShadowHandOpenAI_FF", "ShadowHandOpenAI_LSTM", "AllegroHand", "AllegroHandFF", "AllegroHandLSTM",
         "AllegroKuka", "AllegroKukaLSTM", "AllegroKukaTwoArmsLSTM", "AllegroHandDextremeADR", "AllegroHandDextremeManualDR"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--src", default="/root/reference/isaacgymenvs/cfg")
    ap.add_argument("--out", default=OUT)
    args = ap.parse_args()

    def node(rel):
        with open(os.path.join(args.src, rel + ".yaml")) as f:
            return yaml.safe_load(f) or {}

    table = {"config": node("config")}
    table["config"].pop("hydra", None)      # run-dir plumbing of the trainer, never read by a task
    for t in TASKS:
        table["task/" + t] = node(os.path.join("task", t))
    env_dir = os.path.join(args.src, "task", "env")
    for f in sorted(os.listdir(env_dir)):
        if f.endswith(".yaml"):
            table["task/env/" + f[:-5]] = node(os.path.join("task", "env", f[:-5]))
    with open(args.out, "w") as f:
        json.dump({"source": "isaacgymenvs/cfg (values only; see tools/transcribe_cfg.py)", "nodes": table}, f, indent=1)
        f.write("\n")
    print(f"wrote {os.path.relpath(args.out)}: {len(table)} nodes")


if __name__ == "__main__":
    main()
