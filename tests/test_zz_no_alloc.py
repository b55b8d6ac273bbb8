"""GPU: ``ha_adr_queue_step_static`` -- the allocation-free variant of ``ha_adr_queue_step`` whose scratch (scan state, tagged
per-CTA counts, compacted env list) is a file-scope static ``__device__`` array inside the library instead of three
caller-provided device buffers (``include/handarm_b200.h``).  It has no counterpart in the reference (added
for callers that cannot call cudaMalloc); the check is against the existing entry point on the same inputs.

Collected last on purpose (``test_zz_``).  Self-contained: passes alone in a fresh process
and after the other tests (the entry point clears what must start at zero on every call).

Tolerance: bit-exact, like ``tests/test_adr_update_gpu.py`` -- both entry points run the same two kernels (integer scan, ring
writes in list order, float64 mean in a fixed order)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


class _Side:
    """Persistent outputs of one ADR queue set (rings, append counters, summary) and, for the existing entry point, its scratch."""

    def __init__(self, lib_mod, n, n_queues, cap, worker_types, adr_modes, with_scratch):
        C = lib_mod.CONSTS
        i64 = dict(dtype=torch.int64, device=DEV)
        self.queues = torch.zeros(n_queues, cap, dtype=torch.float32, device=DEV)
        self.appended = torch.zeros(n_queues, **i64)
        self.summary = torch.full((n_queues, 2), -1.0, dtype=torch.float64, device=DEV)
        b = lib_mod.HaAdrQueueBuffers()
        b.worker_types, b.adr_modes = worker_types.data_ptr(), adr_modes.data_ptr()
        b.queues, b.appended, b.summary = self.queues.data_ptr(), self.appended.data_ptr(), self.summary.data_ptr()
        if with_scratch:
            self.list = torch.zeros(n, dtype=torch.int32, device=DEV)
            self.state = torch.zeros(4, **i64)
            self.block_counts = torch.zeros(C["HA_OBSERVER_MAX_CTAS"], **i64)
            b.list, b.state, b.block_counts = self.list.data_ptr(), self.state.data_ptr(), self.block_counts.data_ptr()
        else:
            b.list, b.state, b.block_counts = None, None, None      # ignored by the static variant
        self.b = b


def _inputs(n, n_queues, seed):
    g = torch.Generator().manual_seed(seed)
    reset = (torch.rand(n, generator=g) < 0.3).long().to(DEV)
    applied = (torch.rand(n, generator=g) < 0.1).long().to(DEV)
    objective = torch.randint(0, 40, (n,), generator=g).float().to(DEV)
    return reset, applied, objective


@pytest.mark.parametrize("n,n_queues,cap", [(20000, 24, 8), (65536, 10, 256), (77, 4, 16)])
def test_static_scratch_entry_point_matches_existing(n, n_queues, cap):
    from handarm_b200 import _lib
    lib = _lib.load()
    g = torch.Generator().manual_seed(n)
    worker_types = (torch.rand(n, generator=g) < 0.6).long().to(DEV)
    adr_modes = torch.randint(0, n_queues, (n,), generator=g).to(DEV)
    ref = _Side(_lib, n, n_queues, cap, worker_types, adr_modes, with_scratch=True)
    new = _Side(_lib, n, n_queues, cap, worker_types, adr_modes, with_scratch=False)
    stream = ctypes.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)

    def call(fn, side, reset, applied, objective, all_envs=0):
        side.b.reset_buf, side.b.apply_reset_buf, side.b.objective = reset.data_ptr(), applied.data_ptr(), objective.data_ptr()
        _lib.check(fn(n, n_queues, cap, all_envs, ctypes.byref(side.b), stream))

    # warm-up of both entry points (lazy module loading happens here, outside of the free-memory bracket below)
    warm = _inputs(n, n_queues, 1)
    call(lib.ha_adr_queue_step, ref, *warm)
    call(lib.ha_adr_queue_step_static, new, *warm)
    torch.cuda.synchronize()
    assert torch.equal(new.queues, ref.queues) and torch.equal(new.appended, ref.appended) and torch.equal(new.summary, ref.summary)

    steps = [_inputs(n, n_queues, 10 + k) for k in range(3)]
    torch.cuda.synchronize()
    free_before, _ = torch.cuda.mem_get_info(DEV)
    for k, inp in enumerate(steps):          # calls in a row: the shared array keeps the previous call's contents
        call(lib.ha_adr_queue_step_static, new, *inp, all_envs=int(k == 2))
    torch.cuda.synchronize()
    free_after, _ = torch.cuda.mem_get_info(DEV)
    assert free_after == free_before, "ha_adr_queue_step_static changed the device's free memory (it must not allocate)"
    for k, inp in enumerate(steps):
        call(lib.ha_adr_queue_step, ref, *inp, all_envs=int(k == 2))
    torch.cuda.synchronize()
    assert int(ref.appended.sum()) > 0                                  # the case does append something
    assert torch.equal(new.appended, ref.appended)
    assert torch.equal(new.queues, ref.queues)
    assert torch.equal(new.summary, ref.summary)


def test_static_scratch_two_calls_checked_one_by_one():
    """Two calls in a row, each compared with the existing entry point right after it ran."""
    from handarm_b200 import _lib
    lib = _lib.load()
    n, n_queues, cap = 12345, 16, 32
    g = torch.Generator().manual_seed(3)
    worker_types = (torch.rand(n, generator=g) < 0.5).long().to(DEV)
    adr_modes = torch.randint(0, n_queues, (n,), generator=g).to(DEV)
    ref = _Side(_lib, n, n_queues, cap, worker_types, adr_modes, with_scratch=True)
    new = _Side(_lib, n, n_queues, cap, worker_types, adr_modes, with_scratch=False)
    stream = ctypes.c_void_p(torch.cuda.current_stream(DEV).cuda_stream)
    for k in range(2):
        reset, applied, objective = _inputs(n, n_queues, 100 + k)
        for fn, side in ((lib.ha_adr_queue_step_static, new), (lib.ha_adr_queue_step, ref)):
            side.b.reset_buf, side.b.apply_reset_buf, side.b.objective = reset.data_ptr(), applied.data_ptr(), objective.data_ptr()
            _lib.check(fn(n, n_queues, cap, 0, ctypes.byref(side.b), stream))
        torch.cuda.synchronize()
        assert torch.equal(new.appended, ref.appended), k
        assert torch.equal(new.queues, ref.queues), k
        assert torch.equal(new.summary, ref.summary), k
        assert float(new.summary[:, 1].sum()) > 0


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    from handarm_b200 import _lib
    lib = _lib.load()
    limit = _lib.CONSTS["HA_ADR_STATIC_MAX_ENVS"]
    n = limit + 1
    worker_types = torch.zeros(n, dtype=torch.int64, device=DEV)
    adr_modes = torch.zeros(n, dtype=torch.int64, device=DEV)
    side = _Side(_lib, n, 2, 8, worker_types, adr_modes, with_scratch=False)
    reset, applied, objective = _inputs(n, 2, 0)
    side.b.reset_buf, side.b.apply_reset_buf, side.b.objective = reset.data_ptr(), applied.data_ptr(), objective.data_ptr()
    rc = lib.ha_adr_queue_step_static(n, 2, 8, 0, ctypes.byref(side.b), ctypes.c_void_p(torch.cuda.current_stream(DEV).cuda_stream))
    assert rc == _lib.CONSTS["HA_ERR_INVALID"]
    msg = lib.ha_last_error().decode()
    assert "HA_ADR_STATIC_MAX_ENVS" in msg and str(limit) in msg
    torch.cuda.synchronize()
    assert int(side.appended.sum()) == 0 and float(side.summary.min()) == -1.0      # nothing ran
