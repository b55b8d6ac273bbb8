"""CPU: the build-time snapshot of the reference modules (oracle/_ref, git-ignored, made by oracle/make_ref_snapshot.py from
__graft_entry__.build()) must be self-sufficient: `bench.py --impl reference` runs on a box without /root/reference and imports
every reference module the loader names from the snapshot alone.  A stale snapshot (a module added to the loader after the last
build) fails here instead of on the GPU box.
"""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SNAP = os.path.join(ROOT, "oracle", "_ref")


@pytest.mark.skipif(not os.path.isdir(SNAP), reason="no oracle/_ref snapshot (run __graft_entry__.build() where /root/reference exists)")
def test_snapshot_loads_every_reference_module_alone():
    code = (
        "import os, sys\n"
        "from oracle import reference_loader as rl\n"
        "assert os.path.realpath(rl.REFERENCE_ROOT) == os.path.realpath(sys.argv[1]), rl.REFERENCE_ROOT\n"
        "mods = rl.load_reference(); rl.load_rlgames_utils()\n"
        "root = os.path.realpath(sys.argv[1]) + os.sep\n"
        "for name, m in mods.items():\n"
        "    f = (getattr(m, '__file__', '') or '').replace('<patched>', '')\n"
        "    assert os.path.realpath(f).startswith(root), (name, f)\n"
        "print(len(mods))\n"
    )
    env = dict(os.environ, HANDARM_REFERENCE_ROOT=SNAP)
    r = subprocess.run([sys.executable, "-c", code, SNAP], cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    assert int(r.stdout.strip().splitlines()[-1]) >= 4
