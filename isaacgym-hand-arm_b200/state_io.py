"""AllegroKuka initial-state files (``allegro_kuka_base.py:1506-1592``): the byte format ``dump_env_states`` appends and
``load_initial_states`` reads, so states recorded from a real PhysX run can be replayed as inputs of the synthetic simulator.

File = concatenation of chunks; chunk = ``u32be num_states`` ``u32be len`` ``torch.save(root_states [S, A, 13])`` ``u32be len``
``torch.save(dof_states [S, D, 2])``.  Host-side I/O only (no kernel): the tensors are handed to ``SyntheticSim`` / the reset path."""
from __future__ import annotations

import io
from typing import BinaryIO, Iterable, Tuple

import torch


def _write_tensor(t: torch.Tensor, stream: BinaryIO) -> None:
    buf = io.BytesIO()
    torch.save(t, buf)
    data = buf.getbuffer()
    stream.write(int(len(data)).to_bytes(4, "big"))               # (:1511-1512)
    stream.write(data)


def append_states(path: str, chunks: Iterable[Tuple[torch.Tensor, torch.Tensor]]) -> int:
    """Append ``(root_states [S, A, 13], dof_states [S, D, 2])`` chunks; returns the bytes written (``dump_env_states`` :1514-1543)."""
    out = io.BytesIO()
    for root_states, dof_states in chunks:
        assert root_states.shape[0] == dof_states.shape[0]
        out.write(int(root_states.shape[0]).to_bytes(4, "big"))
        _write_tensor(root_states.detach().cpu(), out)
        _write_tensor(dof_states.detach().cpu(), out)
    data = out.getbuffer()
    if data.nbytes:
        with open(path, "ab") as f:
            f.write(data)
    return data.nbytes


def load_initial_states(path: str) -> Tuple[torch.Tensor, torch.Tensor]:
    """``(initial_root_state_tensors [T, A, 13], initial_dof_state_tensors [T, D, 2])`` (``load_initial_states`` :1545-1592).
    A truncated trailing chunk ends the file like the reference's ``break`` does; an empty file raises."""
    roots, dofs = [], []
    with open(path, "rb") as f:
        def read(n):
            b = f.read(n)
            if len(b) < n:
                raise EOFError
            return b
        while True:
            try:
                int.from_bytes(read(4), "big")                   # num_states of the chunk (the tensors carry it too)
                root_bytes = read(int.from_bytes(read(4), "big"))
                dof_bytes = read(int.from_bytes(read(4), "big"))
            except EOFError:
                break
            roots.append(torch.load(io.BytesIO(root_bytes)))
            dofs.append(torch.load(io.BytesIO(dof_bytes)))
    if not roots:
        raise RuntimeError(f"no states in {path}")
    root, dof = torch.cat(roots), torch.cat(dofs)
    assert root.shape[0] == dof.shape[0]
    return root, dof
