"""ORACLE INFRASTRUCTURE (tests / bench CPU arm only) -- snapshot of the reference's hot-path modules into ``oracle/_ref/``.

``/root/reference`` exists only in the build container.  The GPU box gets a copy of this repository (git-ignored build outputs
included), so the few UNMODIFIED reference files the loader imports (the task classes of the five configs, their
base class and the jit helpers) are copied -- by this script, at build time, never by hand and never into git history
(``oracle/_ref/`` is git-ignored) -- to ``oracle/_ref/isaacgymenvs/...`` keeping their relative paths.  ``reference_loader`` falls
back to that directory when ``/root/reference`` is absent, so ``bench.py --impl reference`` times the reference's OWN
``pre_physics_step`` / ``post_physics_step`` on the box's host cores and the oracle-vs-reference tests can run there too.

The file list is not written down: it is whatever the loader actually imports from the reference tree (``sys.modules`` entries
whose ``__file__`` lies under it), plus the one file the loader reads as text.  Nothing in the product imports from ``oracle/``.

Usage: ``python -m oracle.make_ref_snapshot`` (``__graft_entry__.build()`` runs it when ``/root/reference`` is present).
"""
from __future__ import annotations

import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
DEST = os.path.join(HERE, "_ref")
SOURCE = "/root/reference"


def snapshot(source: str = SOURCE, dest: str = DEST) -> list:
    if not os.path.isdir(os.path.join(source, "isaacgymenvs", "tasks")):
        raise RuntimeError(f"no reference tree under {source}")
    os.environ["HANDARM_REFERENCE_ROOT"] = source
    sys.path.insert(0, os.path.dirname(HERE))
    from oracle import reference_loader as rl
    assert os.path.realpath(rl.REFERENCE_ROOT) == os.path.realpath(source), "reference_loader was imported with another root"
    rl.load_reference()
    rl.load_rlgames_utils()
    files = set()
    root = os.path.realpath(source) + os.sep
    for mod in list(sys.modules.values()):
        path = getattr(mod, "__file__", None) or ""
        path = path.replace("<patched>", "")
        if path and os.path.realpath(path).startswith(root) and path.endswith(".py"):
            files.add(os.path.realpath(path))
    for rel in ("isaacgymenvs/utils/rna_util.py", "isaacgymenvs/tasks/dextreme/allegro_hand_dextreme.py", "LICENSE.txt"):
        p = os.path.join(source, rel)
        if os.path.exists(p):
            files.add(os.path.realpath(p))
    if os.path.isdir(dest):
        shutil.rmtree(dest)
    copied = []
    for path in sorted(files):
        rel = os.path.relpath(path, os.path.realpath(source))
        out = os.path.join(dest, rel)
        os.makedirs(os.path.dirname(out), exist_ok=True)
        shutil.copyfile(path, out)
        copied.append(rel)
    with open(os.path.join(dest, "SNAPSHOT.txt"), "w") as f:
        f.write("Unmodified files of maltemosbach/isaacgym-hand-arm copied by oracle/make_ref_snapshot.py (test infrastructure; git-ignored).\n")
        f.write("\n".join(copied) + "\n")
    return copied


if __name__ == "__main__":
    for rel in snapshot():
        print(rel)
