"""CPU: the C-ABI shared library loads without a GPU, exports every symbol ``include/handarm_b200.h`` declares,
agrees with the header on struct sizes, and reproduces the SURVEY.md 8d byte accounting.  No compute calls."""
import ctypes
import os

import pytest

from handarm_b200 import _lib


@pytest.fixture(scope="module")
def lib():
    if not os.path.exists(_lib.LIB_PATH):
        import __graft_entry__ as g
        g.build()
    return _lib.load()


def test_exports_every_declared_symbol(lib):
    raw = ctypes.CDLL(_lib.LIB_PATH)
    names = _lib.exported_functions()
    assert {"ha_kuka_pre_step", "ha_kuka_post_step", "ha_kuka_observe", "ha_kuka_reset_idx", "ha_last_error", "ha_abi_version"} <= set(names)
    missing = [f for f in names if not hasattr(raw, f)]
    assert not missing, missing


def test_abi_version_and_struct_sizes(lib):
    assert lib.ha_abi_version() == _lib.CONSTS["HA_ABI_VERSION"]
    assert ctypes.sizeof(_lib.HaKukaParams) == lib.ha_sizeof_kuka_params()
    assert ctypes.sizeof(_lib.HaKukaBuffers) == lib.ha_sizeof_kuka_buffers()
    assert ctypes.sizeof(_lib.HaKukaDynamic) == lib.ha_sizeof_kuka_dynamic() == 32


def test_obs_dims_match_reference_layouts(lib):
    # allegro_kuka_base.py:185-220 (single arm), allegro_kuka_two_arms.py:176-211 (two arms)
    assert lib.ha_kuka_obs_dim(1, 4) == 117
    assert lib.ha_kuka_obs_dim(1, 1) == 99
    assert lib.ha_kuka_obs_dim(2, 1) == 169
    assert lib.ha_kuka_obs_dim(2, 4) == 196
    assert lib.ha_kuka_obs_dim(3, 4) == -1


def _params(n_arms=1, K=4):
    p = _lib.HaKukaParams()
    p.abi_version = _lib.CONSTS["HA_ABI_VERSION"]
    p.n_envs, p.n_arms, p.n_dofs, p.dofs_per_env = 32768, n_arms, 23 * n_arms, 23 * n_arms
    p.n_bodies, p.n_tips, p.n_keypoints = 27 if n_arms == 1 else 51, 4 * n_arms, K
    p.obs_dim = _lib.load().ha_kuka_obs_dim(n_arms, K)
    p.palm_body[0], p.object_body, p.success_steps = 7, 24, 1
    for i, t in enumerate((11, 15, 19, 23)):
        p.tip_body[i] = t
    return p


def test_algorithmic_bytes_per_env_step(lib):
    """cfg 3 = 609 read + 613 written by K_post, 504 by K_pre = 1726 B / env-step."""
    p = _params()
    assert lib.ha_kuka_post_bytes_per_env_step(ctypes.byref(p)) == 609 + 613
    assert lib.ha_kuka_pre_bytes_per_env_step(ctypes.byref(p)) == 504
    assert lib.ha_kuka_bytes_per_env_step(ctypes.byref(p)) == 1726
    p2 = _params(n_arms=2, K=1)          # cfg 4: 901 + 805 + 964 = 2670
    assert lib.ha_kuka_bytes_per_env_step(ctypes.byref(p2)) == 2670


def test_hand_algorithmic_bytes_per_env_step(lib):
    """ShadowHand full 640 + 660 + 416 = 1716, full_state 856 + 876 + 416 = 2148, AllegroHand full_state 364 + 384 + 336 = 1084."""
    C = _lib.CONSTS

    def hand(shadow, D, A, T, obs_type, obs_dim):
        p = _lib.HaHandParams()
        p.is_shadow, p.n_dofs, p.n_actions, p.n_tips, p.obs_type, p.obs_dim, p.states_dim = shadow, D, A, T, obs_type, obs_dim, 0
        return p

    sh = hand(1, 24, 20, 5, C["HA_HAND_OBS_FULL"], 157)
    assert lib.ha_hand_post_bytes_per_env_step(ctypes.byref(sh)) == 640 + 660
    assert lib.ha_hand_pre_bytes_per_env_step(ctypes.byref(sh)) == 416
    sh = hand(1, 24, 20, 5, C["HA_HAND_OBS_FULL_STATE"], 211)
    assert lib.ha_hand_bytes_per_env_step(ctypes.byref(sh)) == 2148
    ah = hand(0, 16, 16, 0, C["HA_HAND_OBS_FULL_STATE"], 88)
    assert lib.ha_hand_post_bytes_per_env_step(ctypes.byref(ah)) == 364 + 384
    assert lib.ha_hand_bytes_per_env_step(ctypes.byref(ah)) == 1084


def test_stat_accumulator_initial_values(lib):
    C = _lib.CONSTS
    for i in range(C["HA_STATS_LEN"]):
        want = -2 ** 31 if i in (C["HA_KUKA_STAT_MAX_TRUE_OBJECTIVE"], C["HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE"]) else 0
        assert lib.ha_kuka_stat_accum_init(i) == want


def test_invalid_arguments_are_rejected_without_a_gpu(lib):
    """Validation happens before any CUDA call: wrong ABI version / dims / null buffers give HA_ERR_INVALID and a message."""
    p, b = _params(), _lib.HaKukaBuffers()
    p.abi_version = 999
    rc = lib.ha_kuka_post_step(ctypes.byref(p), ctypes.byref(b), None)
    assert rc == _lib.CONSTS["HA_ERR_INVALID"] and b"abi_version" in lib.ha_last_error()
    p = _params()
    p.n_tips = 5
    assert lib.ha_kuka_post_step(ctypes.byref(p), ctypes.byref(b), None) == _lib.CONSTS["HA_ERR_INVALID"]
    p = _params()
    rc = lib.ha_kuka_pre_step(ctypes.byref(p), ctypes.byref(b), None, None, None)
    assert rc == _lib.CONSTS["HA_ERR_INVALID"] and b"null" in lib.ha_last_error()
    with pytest.raises(_lib.HandArmError):
        _lib.check(rc)


def test_product_has_no_cpu_fallback():
    """A task built on the CPU device has no engine: the hot path raises instead of computing on the host."""
    import torch
    import handarm_b200.config as hcfg
    from handarm_b200.sim import KUKA_LAYOUT, SyntheticSim
    from handarm_b200.tasks.allegro_kuka import AllegroKukaReorientation
    cfg = hcfg.load_task_config("AllegroKuka", num_envs=8, sim_device="cpu")
    cfg["sim"]["use_gpu_pipeline"] = False
    task = AllegroKukaReorientation(cfg, "cpu", "cpu", 0, True, sim=SyntheticSim(8, KUKA_LAYOUT, "cpu"))
    assert task.num_obs == 117 and task.obs_buf.shape == (8, 117) and task.reset_buf.dtype == torch.int64 and int(task.reset_buf.sum()) == 8
    with pytest.raises(_lib.HandArmLibraryError):
        task.post_physics_step()
    with pytest.raises(_lib.HandArmLibraryError):
        task.pre_physics_step(torch.zeros(8, 23))
