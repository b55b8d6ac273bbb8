"""CPU: the AllegroKuka initial-state file format both ways against the reference's own methods
(``allegro_kuka_base.py:1506-1592``): our loader reads what ``dump_env_states`` wrote, ``load_initial_states`` reads what we wrote."""
import random

import pytest
import torch

from oracle import reference_loader as rl

pytestmark = [pytest.mark.reference, pytest.mark.skipif(not rl.reference_available(), reason="needs /root/reference")]


def _bare(tmp, n_envs=3):
    akb = rl.load_reference()["akb"]
    cls = type("BareKuka", (akb.AllegroKukaBase,), {})
    cls.__abstractmethods__ = frozenset()
    t = object.__new__(cls)
    t.save_states_filename = str(tmp / "ref_states.bin")
    t.load_states_filename = str(tmp / "our_states.bin")
    return t


def test_state_files_round_trip_with_the_reference(tmp_path):
    from handarm_b200 import state_io
    g = torch.Generator().manual_seed(0)
    ref = _bare(tmp_path)
    # reference writes: 2 envs with 40 / 25 recorded steps -> min(len // 10, 50) sampled states each
    ref.episode_root_state_tensors = [[torch.randn(4, 13, generator=g) for _ in range(40)], [torch.randn(4, 13, generator=g) for _ in range(25)], []]
    ref.episode_dof_states = [[torch.randn(23, 2, generator=g) for _ in range(40)], [torch.randn(23, 2, generator=g) for _ in range(25)], []]
    pool_root = [torch.stack(x) for x in ref.episode_root_state_tensors[:2]]
    random.seed(3)
    ref.dump_env_states([0, 1, 2])
    root, dof = state_io.load_initial_states(ref.save_states_filename)
    assert root.shape == (4 + 2, 4, 13) and dof.shape == (6, 23, 2)
    for i in range(4):          # every loaded state is one of the recorded ones of env 0
        assert (pool_root[0] == root[i]).all(dim=-1).all(dim=-1).any()
    # we write, the reference reads
    chunks = [(torch.randn(5, 4, 13, generator=g), torch.randn(5, 23, 2, generator=g)), (torch.randn(2, 4, 13, generator=g), torch.randn(2, 23, 2, generator=g))]
    state_io.append_states(ref.load_states_filename, chunks[:1])
    state_io.append_states(ref.load_states_filename, chunks[1:])
    try:
        ref.load_initial_states()
    except UnboundLocalError:
        pytest.skip("the reference's reader re-parses stale buffers in its finally block at end of file on this Python")
    assert ref.num_initial_states >= 7
    assert torch.equal(ref.initial_root_state_tensors[:7], torch.cat([c[0] for c in chunks]))
    assert torch.equal(ref.initial_dof_state_tensors[:7], torch.cat([c[1] for c in chunks]))
    r2, d2 = state_io.load_initial_states(ref.load_states_filename)
    assert torch.equal(r2, torch.cat([c[0] for c in chunks])) and torch.equal(d2, torch.cat([c[1] for c in chunks]))
