"""ORACLE INFRASTRUCTURE (tests only) -- runs the UNMODIFIED reference classes from /root/reference.

The reference's task modules import ``isaacgym`` (closed binary), ``gym``, ``hydra`` ... none of which
exist here.  This loader installs stub modules and namespace-package shims so
that ``isaacgymenvs/tasks/allegro_kuka/*.py`` import *from where they lie*, then builds task objects
with ``object.__new__`` + attribute injection and a recording fake ``gym``.  The reference's own
``pre_physics_step`` / ``post_physics_step`` then run on CPU torch on our synthetic state.

Usable where ``/root/reference`` exists (the build container) or where ``oracle/_ref/`` holds the snapshot of the same files
written by ``oracle/make_ref_snapshot.py`` (the GPU box).  Nothing in the product imports it.
"""
from __future__ import annotations

import contextlib
import importlib
import os
import sys
import types
from typing import Dict, List, Optional

import numpy as np
import torch

def _default_root() -> str:
    """``/root/reference`` in the build container; on the GPU box (where it does not exist) the snapshot of the same unmodified
    files that ``oracle/make_ref_snapshot.py`` wrote to ``oracle/_ref/`` at build time (git-ignored, travels with the repo copy)."""
    if os.path.isdir("/root/reference/isaacgymenvs/tasks"):
        return "/root/reference"
    return os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


REFERENCE_ROOT = os.environ.get("HANDARM_REFERENCE_ROOT") or _default_root()


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "isaacgymenvs", "tasks"))


class _Any:
    """Swallows any attribute access / call; refuses dunders (inspect needs that)."""

    def __init__(self, *a, **k):
        pass

    def __call__(self, *a, **k):
        return _Any()

    def __getattr__(self, n):
        if n.startswith("__"):
            raise AttributeError(n)
        return _Any()


def _stub(name: str, **kw) -> types.ModuleType:
    m = types.ModuleType(name)
    m.__dict__.update(kw)
    sys.modules[name] = m
    return m


_loaded: Dict[str, types.ModuleType] = {}


def load_reference() -> Dict[str, types.ModuleType]:
    """Import the reference's hot-path modules; returns {short name: module}."""
    if _loaded:
        return _loaded
    if not reference_available():
        raise RuntimeError(f"reference not found under {REFERENCE_ROOT}")

    def lazy(n):
        if n.startswith("__"):
            raise AttributeError(n)
        return _Any()

    ig = _stub("isaacgym")
    for sub in ("gymapi", "gymtorch", "gymutil", "torch_utils"):
        m = _stub("isaacgym." + sub)
        m.__getattr__ = lazy
        setattr(ig, sub, m)
    gt = sys.modules["isaacgym.gymtorch"]
    gt.wrap_tensor = lambda t: t
    gt.unwrap_tensor = lambda t: t
    if "gym" not in sys.modules:
        g = _stub("gym")
        g.spaces = _stub("gym.spaces", Box=_Any, Dict=_Any)
        g.Space = object
    if "omegaconf" not in sys.modules:
        _stub("omegaconf", DictConfig=dict, ListConfig=list, OmegaConf=_Any)
    if "tkinter" not in sys.modules:
        _stub("tkinter", W="w")
    root = os.path.join(REFERENCE_ROOT, "isaacgymenvs")
    for name, sub in [("isaacgymenvs", ""), ("isaacgymenvs.tasks", "/tasks"), ("isaacgymenvs.utils", "/utils"),
                      ("isaacgymenvs.tasks.base", "/tasks/base"), ("isaacgymenvs.tasks.allegro_kuka", "/tasks/allegro_kuka"),
                      ("isaacgymenvs.tasks.dextreme", "/tasks/dextreme")]:
        _stub(name).__path__ = [root + sub]
    if not hasattr(np, "Inf"):
        np.Inf = np.inf
    _loaded["tju"] = importlib.import_module("isaacgymenvs.utils.torch_jit_utils")
    _loaded["akb"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_base")
    _loaded["akr"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_reorientation")
    _loaded["aku"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_utils")
    _loaded["akg"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_regrasping")
    _loaded["akt"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_throw")
    _loaded["vt"] = importlib.import_module("isaacgymenvs.tasks.base.vec_task")
    _loaded["ak2"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_two_arms")
    _loaded["ak2g"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_two_arms_regrasping")
    _loaded["ak2r"] = importlib.import_module("isaacgymenvs.tasks.allegro_kuka.allegro_kuka_two_arms_reorientation")
    _loaded["sh"] = importlib.import_module("isaacgymenvs.tasks.shadow_hand")
    _loaded["ah"] = importlib.import_module("isaacgymenvs.tasks.allegro_hand")
    _loaded["adrv"] = importlib.import_module("isaacgymenvs.tasks.dextreme.adr_vec_task")
    _loaded["dx"] = _load_dextreme(root)
    return _loaded


def _load_dextreme(root: str) -> types.ModuleType:
    """``allegro_hand_dextreme.py`` executed from where it lies with TWO source-level patches, both needed only to run it on
    CPU torch 2.x: ``:1630`` passes an int64 tensor as the condition of ``torch.where``, which torch >= 2 rejects (``!= 0`` is the
    meaning old torch gave it); ``:969`` shifts the action queue with an overlapping slice assignment, which the CPU
    backend refuses ("refer to a single memory location") -- ``.clone()`` of the source gives the shift the CUDA backend performs."""
    path = os.path.join(root, "tasks", "dextreme", "allegro_hand_dextreme.py")
    src = open(path).read()
    old, new = "torch.where(goal_reached, hold_count_buf + 1", "torch.where(goal_reached != 0, hold_count_buf + 1"
    assert src.count(old) == 1, "the reference changed: the documented torch.where patch no longer applies"
    name = "isaacgymenvs.tasks.dextreme.allegro_hand_dextreme"
    mod = types.ModuleType(name)
    mod.__file__ = path
    sys.modules[name] = mod
    old2, new2 = "self.prev_actions_queue[:, :-1].detach()", "self.prev_actions_queue[:, :-1].clone()"
    assert src.count(old2) == 1, "the reference changed: the documented queue-shift patch no longer applies"
    patched = src.replace(old, new).replace(old2, new2)
    shown = path + "<patched>"                                # torch.jit.script re-reads the source through linecache
    import linecache
    linecache.cache[shown] = (len(patched), None, patched.splitlines(True), shown)
    mod.__file__ = shown
    exec(compile(patched, shown, "exec"), mod.__dict__)
    return mod


class FakeGym:
    """Records every call the task makes into the simulator and returns None."""

    def __init__(self):
        self.calls: List[tuple] = []

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)

        def rec(*args, **kwargs):
            self.calls.append((name, args))
            return None

        return rec


KUKA_EPISODE_KEYS = ["raw_fingertip_delta_rew", "raw_hand_delta_penalty", "raw_lifting_rew", "raw_keypoint_rew", "fingertip_delta_rew",
                     "hand_delta_penalty", "lifting_rew", "lift_bonus_rew", "keypoint_rew", "bonus_rew", "kuka_actions_penalty",
                     "allegro_actions_penalty"]


def make_reference_kuka(st: Dict[str, torch.Tensor], c, device: str = "cpu") -> object:
    """An ``AllegroKukaReorientation`` instance (reference class) aliasing the tensors of ``st``.

    ``st`` is the dict layout of ``oracle/kuka_state.py``; ``c`` a ``KukaConsts``.  The instance owns
    views of the same storage wherever the reference keeps views (``arm_hand_dof_pos`` ...).
    ``device`` other than "cpu" moves the state there first (the reference's eager GPU path, a context number of ``bench.py``);
    ``st`` then no longer aliases the instance's tensors.
    """
    mods = load_reference()
    if device != "cpu":
        st = {k: (v.to(device) if torch.is_tensor(v) else v) for k, v in st.items()}
    subtask = getattr(c, "subtask", "reorientation")
    cls = {"reorientation": lambda: mods["akr"].AllegroKukaReorientation, "regrasping": lambda: mods["akg"].AllegroKukaRegrasping,
           "throw": lambda: mods["akt"].AllegroKukaThrow}[subtask]()
    env = object.__new__(cls)
    N = st["reset_buf"].shape[0]
    K = st["object_keypoint_offsets"].shape[1]
    e = env.__dict__
    e.update(
        gym=FakeGym(), sim=None, device=device, rl_device=device, num_environments=N, viewer=None, debug_viz=False, headless=True,
        num_arm_dofs=c.n_arm_dofs, num_hand_dofs=c.n_dofs - c.n_arm_dofs, num_hand_arm_dofs=c.n_dofs, num_finger_dofs=4,
        num_allegro_fingertips=c.n_tips, num_keypoints=K, obs_type="full_state", with_fingertip_force_sensors=False,
        with_dof_force_sensors=False, num_observations=st["obs_buf"].shape[1], privileged_actions=False, use_relative_control=False,
        randomize=False, should_load_initial_states=False, save_states=False, eval_stats=False,
        dt=c.dt, act_moving_average=c.act_moving_average, hand_dof_speed_scale=c.hand_dof_speed_scale, force_scale=c.force_scale,
        force_decay_interval=c.force_decay_interval, clamp_abs_observations=c.clamp_abs_observations,
        max_episode_length=c.max_episode_length, max_consecutive_successes=c.max_consecutive_successes, success_steps=c.success_steps,
        success_tolerance=c.success_tolerance, initial_tolerance=c.initial_tolerance, target_tolerance=c.target_tolerance,
        tolerance_curriculum_increment=0.9, tolerance_curriculum_interval=10 ** 9, keypoint_scale=c.keypoint_scale,
        lifting_bonus_threshold=c.lifting_bonus_threshold, lifting_bonus=c.lifting_bonus, lifting_rew_scale=c.lifting_rew_scale,
        keypoint_rew_scale=c.keypoint_rew_scale, distance_delta_rew_scale=c.distance_delta_rew_scale, reach_goal_bonus=c.reach_goal_bonus,
        kuka_actions_penalty_scale=c.kuka_actions_penalty_scale, allegro_actions_penalty_scale=c.allegro_actions_penalty_scale,
        reset_position_noise_x=c.reset_position_noise[0], reset_position_noise_y=c.reset_position_noise[1],
        reset_position_noise_z=c.reset_position_noise[2], reset_dof_pos_noise_arm=c.reset_dof_pos_noise_arm,
        reset_dof_pos_noise_fingers=c.reset_dof_pos_noise_fingers, reset_dof_vel_noise=c.reset_dof_vel_noise,
        frame_since_restart=0, last_curriculum_update=0, control_freq_inv=1, extras={}, total_successes=0, total_resets=0,
        set_actor_root_state_object_indices=[], randomization_params={},
    )
    env.full_state_size = st["obs_buf"].shape[1]
    env.force_decay = torch.tensor(c.force_decay, dtype=torch.float32, device=device)
    env.force_prob_range = torch.tensor(c.force_prob_range, dtype=torch.float32, device=device)
    env.palm_offset = np.array(c.palm_offset, dtype=np.float32)
    env.fingertip_offsets = np.array(c.tip_offsets, dtype=np.float32)
    env.target_volume_origin = torch.tensor(c.target_volume_origin, dtype=torch.float32, device=device)
    env.target_volume_extent = torch.tensor(c.target_volume_extent, dtype=torch.float32, device=device)
    env.allegro_palm_handle = c.palm_body
    env.allegro_fingertip_handles = torch.tensor(list(c.tip_bodies), dtype=torch.long, device=device)
    env.object_rb_handles = torch.tensor([c.object_body], dtype=torch.long, device=device)
    env.object_rb_masses = torch.tensor([c.object_mass], dtype=torch.float32, device=device)
    # simulator tensors and the reference's views of them (allegro_kuka_base.py:280-295)
    env.root_state_tensor = st["root_state"]
    env.rigid_body_states = st["rigid_body_state"].view(N, -1, 13)
    env.dof_state = st["dof_state"].view(-1, 2)
    env.arm_hand_dof_state = env.dof_state.view(N, -1, 2)[:, : c.n_dofs]
    env.arm_hand_dof_pos = env.arm_hand_dof_state[..., 0]
    env.arm_hand_dof_vel = env.arm_hand_dof_state[..., 1]
    env.num_bodies = env.rigid_body_states.shape[1]
    if subtask == "throw":
        env.bucket_object_indices = st["goal_object_indices"]          # the bucket takes the goal actor's slot
    for name in ("object_indices", "goal_object_indices", "allegro_hand_indices", "goal_states", "object_init_state",
                 "object_keypoint_offsets", "object_scales", "finger_rew_coeffs", "obs_buf", "rew_buf", "reset_buf", "reset_goal_buf",
                 "progress_buf", "randomize_buf", "successes", "prev_episode_successes", "true_objective",
                 "prev_episode_true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
                 "closest_fingertip_dist", "furthest_hand_dist", "prev_targets", "cur_targets", "rb_forces", "random_force_prob"):
        setattr(env, name, st[name])
    env.arm_hand_dof_lower_limits = st["dof_lower"]
    env.arm_hand_dof_upper_limits = st["dof_upper"]
    env.hand_arm_default_dof_pos = st["dof_default"]
    env.obj_keypoint_pos = torch.zeros(N, K, 3, device=device)
    env.goal_keypoint_pos = torch.zeros(N, K, 3, device=device)
    env.action_torques = torch.zeros_like(st["rb_forces"])
    env.rewards_episode = {k: st["rewards_episode"][i] for i, k in enumerate(KUKA_EPISODE_KEYS)}
    env.timeout_buf = torch.zeros(N, dtype=torch.long, device=device)
    return env


@contextlib.contextmanager
def scripted_randoms(queue: List[torch.Tensor]):
    """Feed the reference pre-drawn raw numbers instead of torch's global generator.

    ``queue`` holds, in the reference's consumption order, raw U(0,1) / N(0,1) tensors of exactly the
    shape each call asks for.  ``torch_rand_float(lo, hi, shape)`` becomes ``(hi - lo) * raw + lo``.
    """
    mods = load_reference()
    q = list(queue)

    def pop(shape):
        t = q.pop(0)
        assert tuple(t.shape) == tuple(shape), f"scripted random shape {tuple(t.shape)} != requested {tuple(shape)}"
        return t.clone()

    def fake_rand_float(lower, upper, shape, device=None):
        return (upper - lower) * pop(shape) + lower

    def fake_rand(*shape, **kw):
        shape = shape[0] if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)) else shape
        return pop(shape)

    saved = {}
    patched = [k for k in ("akb", "akr", "akg", "akt", "sh", "ah", "ak2", "ak2g", "ak2r", "dx") if k in mods and hasattr(mods[k], "torch_rand_float")]
    for key in patched:
        saved[key] = mods[key].torch_rand_float
        mods[key].torch_rand_float = fake_rand_float
    old_rand, old_randn, old_rand_like = torch.rand, torch.randn, torch.rand_like
    torch.rand, torch.randn = fake_rand, fake_rand
    torch.rand_like = lambda t, **kw: pop(t.shape)
    try:
        yield q
    finally:
        torch.rand, torch.randn, torch.rand_like = old_rand, old_randn, old_rand_like
        for key in patched:
            mods[key].torch_rand_float = saved[key]


def kuka_random_queue(st: Dict[str, torch.Tensor], c, rnd: torch.Tensor) -> List[torch.Tensor]:
    """Order of draws of one reorientation ``pre_physics_step`` for table ``rnd``."""
    from . import kuka_oracle as ko
    D = c.n_dofs
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    subtask = getattr(c, "subtask", "reorientation")

    def target_draws(who, slot):
        """the draws of one reset_target_pose(who) in the reference's order"""
        obj = [rnd[who, ko.RND_OBJ_POS:ko.RND_OBJ_POS + 3], rnd[who, ko.RND_OBJ_ROT:ko.RND_OBJ_ROT + 3]]      # reset_object_pose
        if subtask == "reorientation":
            return [rnd[who, slot:slot + 3], rnd[who, slot + 3:slot + 6]]
        if subtask == "regrasping":
            return [rnd[who, slot:slot + 3]] + obj
        return [rnd[who, slot + k:slot + k + 1] for k in range(4)] + obj                                          # throw: 4 x (len, 1)

    q = target_draws(goal_ids, ko.RND_GOAL_POS)
    prob = st["random_force_prob"].clone()
    if len(ids) > 0:
        q += target_draws(ids, ko.RND_RESET_GOAL_POS)
        q += [rnd[ids, ko.RND_OBJ_POS:ko.RND_OBJ_POS + 3], rnd[ids, ko.RND_OBJ_ROT:ko.RND_OBJ_ROT + 3], rnd[ids, ko.RND_FORCE_PROB],
              rnd[ids, ko.RND_DOF_POS:ko.RND_DOF_POS + D], rnd[ids, ko.RND_DOF_POS + D:ko.RND_DOF_POS + 2 * D]]
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        prob[ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, ko.RND_FORCE_PROB] + torch.log(pr[1]))
    if c.force_scale > 0.0:
        q.append(rnd[:, ko.RND_FORCE_TEST])
        hit = (rnd[:, ko.RND_FORCE_TEST] < prob).nonzero(as_tuple=False)
        q.append(rnd[hit.squeeze(-1), ko.RND_FORCE_NORMAL:ko.RND_FORCE_NORMAL + 3].unsqueeze(1))
    return q


def reference_kuka_step(env, c, actions: torch.Tensor, rnd: torch.Tensor, st: Dict[str, torch.Tensor]) -> None:
    """pre_physics_step + post_physics_step + the VecTask.step tail (vec_task.py:424) of the reference."""
    with scripted_randoms(kuka_random_queue(st, c, rnd)) as q:
        env.pre_physics_step(actions)
        assert len(q) == 0, f"{len(q)} scripted random tensors were not consumed"
    env.post_physics_step()
    env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)


def harvest(env, st: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Copy the attributes the reference REBINDS (instead of mutating in place) back into a state dict."""
    out = dict(st)
    for name in ("progress_buf", "successes", "true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
                 "closest_fingertip_dist", "furthest_hand_dist", "timeout_buf", "reset_buf", "reset_goal_buf", "randomize_buf",
                 "prev_episode_successes", "prev_episode_true_objective", "random_force_prob", "obs_buf", "rew_buf"):
        out[name] = getattr(env, name)
    out["actions"] = env.actions
    out["rewards_episode"] = torch.stack([env.rewards_episode[k] for k in KUKA_EPISODE_KEYS])
    cum = env.extras.get("episode_cumulative")
    if cum is not None:
        from .kuka_oracle import STEP_TERMS
        out["episode_cumulative"] = torch.stack([cum[k] for k in STEP_TERMS])
    return out


# ---------------------------------------------------------------------------------------------
# ShadowHand / AllegroHand (shadow_hand.py, allegro_hand.py)
# ---------------------------------------------------------------------------------------------
def make_reference_hand(st: Dict[str, torch.Tensor], c) -> object:
    """A reference ``ShadowHand`` / ``AllegroHand`` instance aliasing the tensors of ``st`` (``oracle/hand_state.py`` layout)."""
    mods = load_reference()
    cls = mods["sh"].ShadowHand if c.kind == "shadow" else mods["ah"].AllegroHand
    env = object.__new__(cls)
    N, D = st["reset_buf"].shape[0], c.n_dofs
    env.__dict__.update(
        gym=FakeGym(), sim=None, device="cpu", rl_device="cpu", num_environments=N, viewer=None, debug_viz=False, headless=True,
        num_actions=c.n_actions, num_shadow_hand_dofs=D, num_fingertips=c.n_tips, obs_type=c.obs_type, asymmetric_obs=c.asymmetric,
        randomize=False, randomization_params={}, object_type="pen" if c.ignore_z_rot else "block", up_axis_idx=c.up_axis_idx,
        dt=c.dt, use_relative_control=c.use_relative_control, shadow_hand_dof_speed_scale=c.dof_speed_scale,
        act_moving_average=c.act_moving_average, force_scale=c.force_scale, force_decay_interval=c.force_decay_interval,
        max_episode_length=c.max_episode_length, dist_reward_scale=c.dist_reward_scale, rot_reward_scale=c.rot_reward_scale,
        rot_eps=c.rot_eps, action_penalty_scale=c.action_penalty_scale, success_tolerance=c.success_tolerance,
        reach_goal_bonus=c.reach_goal_bonus, fall_dist=c.fall_dist, fall_penalty=c.fall_penalty,
        max_consecutive_successes=c.max_consecutive_successes, print_success_stat=False, vel_obs_scale=c.vel_obs_scale,
        force_torque_obs_scale=c.force_torque_obs_scale, reset_position_noise=c.reset_position_noise,
        reset_dof_pos_noise=c.reset_dof_pos_noise, reset_dof_vel_noise=c.reset_dof_vel_noise, extras={}, total_successes=0, total_resets=0,
    )
    env.av_factor = torch.tensor(c.av_factor, dtype=torch.float32)
    env.force_decay = torch.tensor(c.force_decay, dtype=torch.float32)
    env.force_prob_range = torch.tensor(c.force_prob_range, dtype=torch.float32)
    env.goal_displacement_tensor = torch.tensor(c.goal_displacement, dtype=torch.float32)
    env.x_unit_tensor = torch.tensor([1.0, 0.0, 0.0]).repeat(N, 1)
    env.y_unit_tensor = torch.tensor([0.0, 1.0, 0.0]).repeat(N, 1)
    env.z_unit_tensor = torch.tensor([0.0, 0.0, 1.0]).repeat(N, 1)
    env.actuated_dof_indices = torch.tensor(c.actuated, dtype=torch.long)
    env.fingertip_handles = torch.tensor(list(c.tip_bodies), dtype=torch.long)
    env.object_rb_handles = torch.tensor([c.object_body], dtype=torch.long)
    env.object_rb_masses = torch.tensor([c.object_mass], dtype=torch.float32)
    env.root_state_tensor = st["root_state"]
    env.rigid_body_states = st["rigid_body_state"].view(N, -1, 13)
    env.dof_state = st["dof_state"].view(-1, 2)
    env.shadow_hand_dof_state = env.dof_state.view(N, -1, 2)[:, :D]
    env.shadow_hand_dof_pos = env.shadow_hand_dof_state[..., 0]
    env.shadow_hand_dof_vel = env.shadow_hand_dof_state[..., 1]
    env.dof_force_tensor = st["dof_force"]
    env.vec_sensor_tensor = st["sensors"]
    env.shadow_hand_dof_lower_limits = st["dof_lower"]
    env.shadow_hand_dof_upper_limits = st["dof_upper"]
    env.shadow_hand_default_dof_pos = st["dof_default"]
    env.shadow_hand_dof_default_pos = st["dof_default"]
    env.shadow_hand_dof_default_vel = torch.zeros(D)
    env.num_bodies = env.rigid_body_states.shape[1]
    for name in ("hand_indices", "object_indices", "goal_object_indices", "goal_states", "goal_init_state", "object_init_state", "obs_buf",
                 "states_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "successes",
                 "consecutive_successes", "prev_targets", "cur_targets", "rb_forces", "random_force_prob"):
        setattr(env, name, st[name])
    env.actions = st["actions"]
    env.timeout_buf = st["timeout_buf"]
    return env


def hand_random_queue(st: Dict[str, torch.Tensor], c, rnd: torch.Tensor) -> List[torch.Tensor]:
    """Order of draws of one hand ``pre_physics_step`` for table ``rnd`` (``HA_HAND_RND_*``)."""
    from . import hand_oracle as ho
    D = c.n_dofs
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)
    q: List[torch.Tensor] = []
    if len(goal_ids) > 0:
        q.append(rnd[goal_ids, ho.RND_GOAL:ho.RND_GOAL + 4])
    prob = st["random_force_prob"].clone()
    if len(ids) > 0:
        q += [rnd[ids, ho.RND_RESET:ho.RND_RESET + 2 * D + 5], rnd[ids, ho.RND_RESET_GOAL:ho.RND_RESET_GOAL + 4], rnd[ids, ho.RND_FORCE_PROB]]
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        prob[ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, ho.RND_FORCE_PROB] + torch.log(pr[1]))
    if c.force_scale > 0.0:
        q.append(rnd[:, ho.RND_FORCE_TEST])
        hit = (rnd[:, ho.RND_FORCE_TEST] < prob).nonzero(as_tuple=False)
        q.append(rnd[hit.squeeze(-1), ho.RND_FORCE_NORMAL:ho.RND_FORCE_NORMAL + 3].unsqueeze(1))
    return q


def reference_hand_step(env, c, actions: torch.Tensor, rnd: torch.Tensor, st: Dict[str, torch.Tensor]) -> None:
    with scripted_randoms(hand_random_queue(st, c, rnd)) as q:
        env.pre_physics_step(actions)
        assert len(q) == 0, f"{len(q)} scripted random tensors were not consumed"
    env.post_physics_step()
    env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)


def harvest_hand(env, st: Dict[str, torch.Tensor], c) -> Dict[str, torch.Tensor]:
    out = dict(st)
    for name in ("progress_buf", "successes", "consecutive_successes", "timeout_buf", "reset_buf", "reset_goal_buf", "randomize_buf",
                 "random_force_prob", "obs_buf", "states_buf", "rew_buf", "rb_forces"):
        out[name] = getattr(env, name)
    out["actions"] = env.actions
    out["obs_out"] = torch.clamp(env.obs_buf, -c.clip_obs, c.clip_obs)
    if c.asymmetric:
        out["states_out"] = torch.clamp(env.states_buf, -c.clip_obs, c.clip_obs)
    return out


# ---------------------------------------------------------------------------------------------
# AllegroKukaTwoArmsRegrasping (allegro_kuka_two_arms.py, allegro_kuka_two_arms_regrasping.py)
# ---------------------------------------------------------------------------------------------
def make_reference_kuka2(st: Dict[str, torch.Tensor], c) -> object:
    """A reference ``AllegroKukaTwoArmsRegrasping`` / ``AllegroKukaTwoArmsReorientation`` instance (``c.subtask``) aliasing the tensors
    of ``st`` (``oracle/kuka2_oracle.make_state``)."""
    from .kuka2_oracle import EPISODE_TERMS
    mods = load_reference()
    cls = (mods["ak2r"].AllegroKukaTwoArmsReorientation if getattr(c, "subtask", "regrasping") == "reorientation"
           else mods["ak2g"].AllegroKukaTwoArmsRegrasping)
    env = object.__new__(cls)
    N = st["reset_buf"].shape[0]
    K = st["object_keypoint_offsets"].shape[1]
    env.__dict__.update(
        gym=FakeGym(), sim=None, device="cpu", rl_device="cpu", num_environments=N, viewer=None, debug_viz=False, headless=True,
        num_arms=c.n_arms, num_arm_dofs=c.n_arm_dofs, num_hand_dofs=16, num_hand_arm_dofs=23, num_finger_dofs=4, num_fingertips=4,
        num_allegro_fingertips=4, num_keypoints=K, obs_type="full_state", num_observations=st["obs_buf"].shape[1],
        privileged_actions=False, use_relative_control=False, randomize=False, should_load_initial_states=False, save_states=False,
        eval_stats=False, dt=c.dt, act_moving_average=c.act_moving_average, force_scale=c.force_scale,
        force_decay_interval=c.force_decay_interval, clamp_abs_observations=c.clamp_abs_observations,
        max_episode_length=c.max_episode_length, max_consecutive_successes=c.max_consecutive_successes, success_steps=c.success_steps,
        success_tolerance=c.success_tolerance, initial_tolerance=c.initial_tolerance, target_tolerance=c.target_tolerance,
        tolerance_curriculum_increment=0.9, tolerance_curriculum_interval=10 ** 9, keypoint_scale=c.keypoint_scale,
        lifting_bonus_threshold=c.lifting_bonus_threshold, lifting_bonus=c.lifting_bonus, lifting_rew_scale=c.lifting_rew_scale,
        keypoint_rew_scale=c.keypoint_rew_scale, distance_delta_rew_scale=c.distance_delta_rew_scale, reach_goal_bonus=c.reach_goal_bonus,
        reset_position_noise_x=c.reset_position_noise[0], reset_position_noise_y=c.reset_position_noise[1],
        reset_position_noise_z=c.reset_position_noise[2], reset_dof_vel_noise=c.reset_dof_vel_noise, frame_since_restart=0,
        last_curriculum_update=0, control_freq_inv=1, extras={}, total_successes=0, total_resets=0,
        set_actor_root_state_object_indices=[], randomization_params={},
    )
    env.full_state_size = st["obs_buf"].shape[1]
    env.force_decay = torch.tensor(c.force_decay, dtype=torch.float32)
    env.force_prob_range = torch.tensor(c.force_prob_range, dtype=torch.float32)
    env.palm_center_offset = torch.tensor(c.palm_offset, dtype=torch.float32).repeat(N, 1)
    env.palm_center_pos = torch.zeros(N, c.n_arms, 3)
    env.fingertip_offsets = torch.tensor(np.asarray(c.tip_offsets, dtype=np.float32)).repeat(N, 1, 1)
    env.target_volume_origin = torch.tensor(c.target_volume_origin, dtype=torch.float32)
    env.target_volume_extent = torch.tensor(c.target_volume_extent, dtype=torch.float32)
    env.allegro_palm_handles = list(c.palm_bodies)
    env.allegro_fingertip_handles = torch.tensor(list(c.tip_bodies), dtype=torch.long)
    env.object_rb_handles = torch.tensor([c.object_body], dtype=torch.long)
    env.object_rb_masses = torch.tensor([c.object_mass], dtype=torch.float32)
    env.root_state_tensor = st["root_state"]
    env.rigid_body_states = st["rigid_body_state"].view(N, -1, 13)
    env.dof_state = st["dof_state"].view(-1, 2)
    env.arm_hand_dof_state = env.dof_state.view(N, -1, 2)[:, : c.n_dofs]
    env.arm_hand_dof_pos = env.arm_hand_dof_state[..., 0]
    env.arm_hand_dof_vel = env.arm_hand_dof_state[..., 1]
    env.num_bodies = env.rigid_body_states.shape[1]
    for name in ("arm_indices", "object_indices", "goal_object_indices", "goal_states", "object_init_state", "object_keypoint_offsets",
                 "object_scales", "obs_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "successes",
                 "prev_episode_successes", "true_objective", "prev_episode_true_objective", "near_goal_steps", "lifted_object",
                 "closest_keypoint_max_dist", "closest_fingertip_dist", "prev_targets", "cur_targets", "rb_forces", "random_force_prob",
                 "pos_noise_coeff"):
        setattr(env, name, st[name])
    env.arm_hand_dof_lower_limits = st["dof_lower"]
    env.arm_hand_dof_upper_limits = st["dof_upper"]
    env.hand_arm_default_dof_pos = st["dof_default"]
    env.obj_keypoint_pos = torch.zeros(N, K, 3)
    env.goal_keypoint_pos = torch.zeros(N, K, 3)
    env.rewards_episode = {k: st["rewards_episode"][i] for i, k in enumerate(EPISODE_TERMS)}
    env.timeout_buf = torch.zeros(N, dtype=torch.long)
    return env


def kuka2_random_queue(st: Dict[str, torch.Tensor], c, rnd: torch.Tensor) -> List[torch.Tensor]:
    """Order of draws of one two-arm regrasping ``pre_physics_step`` for table ``rnd`` (``HA_KUKA2_RND_*``)."""
    from . import kuka2_oracle as k2
    D = c.n_dofs
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    ids = st["reset_buf"].nonzero(as_tuple=False).squeeze(-1)

    def goal_block(rows, base):       # _reset_target: position, side; then reset_object_pose: side, position, rotation (regrasping)
        if getattr(c, "subtask", "regrasping") == "reorientation":            # ... or get_random_quat: uvw (reorientation)
            return [rnd[rows, base:base + 3], rnd[rows, base + 3:base + 4], rnd[rows, base + 4:base + 7]]
        return [rnd[rows, base:base + 3], rnd[rows, base + 3:base + 4], rnd[rows, base + 4:base + 5], rnd[rows, base + 5:base + 8],
                rnd[rows, base + 8:base + 11]]

    q = goal_block(goal_ids, k2.RND_GOAL)
    prob = st["random_force_prob"].clone()
    if len(ids) > 0:
        q += goal_block(ids, k2.RND_RESET_GOAL)
        b = k2.RND_RESET_OBJ
        q += [rnd[ids, b:b + 1], rnd[ids, b + 1:b + 4], rnd[ids, b + 4:b + 7], rnd[ids, k2.RND_FORCE_PROB],
              rnd[ids, k2.RND_DOF_POS:k2.RND_DOF_POS + D], rnd[ids, k2.RND_DOF_POS + D:k2.RND_DOF_POS + 2 * D]]
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        prob[ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, k2.RND_FORCE_PROB] + torch.log(pr[1]))
    if c.force_scale > 0.0:
        q.append(rnd[:, k2.RND_FORCE_TEST])
        hit = (rnd[:, k2.RND_FORCE_TEST] < prob).nonzero(as_tuple=False)
        q.append(rnd[hit.squeeze(-1), k2.RND_FORCE_NORMAL:k2.RND_FORCE_NORMAL + 3].unsqueeze(1))
    return q


def reference_kuka2_step(env, c, actions: torch.Tensor, rnd: torch.Tensor, st: Dict[str, torch.Tensor]) -> None:
    with scripted_randoms(kuka2_random_queue(st, c, rnd)) as q:
        env.pre_physics_step(actions)
        assert len(q) == 0, f"{len(q)} scripted random tensors were not consumed"
    env.post_physics_step()
    env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)


def harvest_kuka2(env, st: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    from .kuka2_oracle import EPISODE_TERMS, STEP_TERMS
    out = dict(st)
    for name in ("progress_buf", "successes", "true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
                 "closest_fingertip_dist", "timeout_buf", "reset_buf", "reset_goal_buf", "randomize_buf", "prev_episode_successes",
                 "prev_episode_true_objective", "random_force_prob", "obs_buf", "rew_buf"):
        out[name] = getattr(env, name)
    out["actions"] = env.actions
    out["rewards_episode"] = torch.stack([env.rewards_episode[k] for k in EPISODE_TERMS])
    cum = env.extras.get("episode_cumulative")
    if cum is not None:
        out["episode_cumulative"] = torch.stack([cum[k] for k in STEP_TERMS])
    return out


# ---------------------------------------------------------------------------------------------
# AllegroHandDextremeADR (tasks/dextreme/allegro_hand_dextreme.py, adr_vec_task.py)
# ---------------------------------------------------------------------------------------------
def make_reference_dextreme(st: Dict[str, torch.Tensor], c) -> object:
    """A reference ``AllegroHandDextremeADR`` instance aliasing the tensors of ``st`` (``oracle/dextreme_state.py`` layout).
    ``apply_randomizations`` (Isaac Gym property DR + host ADR bookkeeping) is replaced by a no-op: the per-env ADR tensors and
    affine parameters of ``st`` play the role of what it sampled."""
    from . import dextreme_oracle as dx
    mods = load_reference()
    env = object.__new__(mods["dx"].AllegroHandDextremeADR)
    N, D = st["reset_buf"].shape[0], c.n_dofs
    env.__dict__.update(
        gym=FakeGym(), sim=None, device="cpu", rl_device="cpu", sim_device="cpu", num_environments=N, viewer=None, debug_viz=False,
        headless=True, num_actions=D, num_hand_dofs=D, num_fingertips=c.n_tips, randomize=True, use_adr=True, randomization_params={},
        up_axis_idx=2, dt=c.dt, control_freq_inv=c.control_freq_inv, use_relative_control=False,
        use_capped_dof_control=c.use_capped_dof_control, max_dof_radians_per_second=c.max_dof_radians_per_second,
        act_moving_average=c.act_moving_average, act_moving_average_scheduled_freq=500, act_moving_average_scheduled_steps=1000000,
        act_moving_average_lower=0.15, act_moving_average_upper=0.2, last_step=0, enable_rna=False, apply_random_quat=c.apply_random_quat,
        random_cube_pose_prob=c.random_cube_pose_prob, enable_random_obs=True, force_scale=c.force_scale,
        force_decay_interval=c.force_decay_interval, max_episode_length=c.max_episode_length, dist_reward_scale=c.dist_reward_scale,
        rot_reward_scale=c.rot_reward_scale, rot_eps=c.rot_eps, action_penalty_scale=c.action_penalty_scale,
        action_delta_penalty_scale=c.action_delta_penalty_scale, success_tolerance=c.success_tolerance,
        reach_goal_bonus=c.reach_goal_bonus, fall_dist=c.fall_dist, fall_penalty=c.fall_penalty,
        max_consecutive_successes=c.max_consecutive_successes, num_success_hold_steps=c.num_success_hold_steps,
        print_success_stat=False, vel_obs_scale=c.vel_obs_scale, force_torque_obs_scale=c.force_torque_obs_scale,
        reset_position_noise=c.reset_position_noise, reset_position_noise_z=c.reset_position_noise_z,
        reset_dof_pos_noise=c.reset_dof_pos_noise, reset_dof_vel_noise=c.reset_dof_vel_noise, extras={}, total_successes=0, total_resets=0,
        clip_obs=c.clip_obs,
    )
    env.apply_randomizations = lambda *a, **k: None
    env.frame = int(st["frame"])
    env.av_factor = torch.tensor(c.av_factor, dtype=torch.float32)
    env.force_decay = torch.tensor(c.force_decay, dtype=torch.float32)
    env.force_prob_range = torch.tensor(c.force_prob_range, dtype=torch.float32)
    env.goal_displacement_tensor = torch.tensor(c.goal_displacement, dtype=torch.float32)
    env.x_unit_tensor = torch.tensor([1.0, 0.0, 0.0]).repeat(N, 1)
    env.y_unit_tensor = torch.tensor([0.0, 1.0, 0.0]).repeat(N, 1)
    env.actuated_dof_indices = torch.arange(D, dtype=torch.long)
    env.fingertip_handles = torch.tensor(list(c.tip_bodies), dtype=torch.long)
    env.palm_link_handle = (c.palm_body,)                                   # the reference stores a 1-tuple (:390, trailing comma)
    env.object_rb_handles = torch.tensor([c.object_body], dtype=torch.long)
    env.object_rb_masses = torch.tensor([c.object_mass], dtype=torch.float32)
    env.all_env_ids = torch.arange(N, dtype=torch.long)
    env.root_state_tensor = st["root_state"]
    env.rigid_body_states = st["rigid_body_state"].view(N, -1, 13)
    env.dof_state = st["dof_state"].view(N, -1, 2)[:, :D]
    env.dof_pos = env.dof_state[..., 0]
    env.dof_vel = env.dof_state[..., 1]
    env.dof_force_tensor = st["dof_force"]
    env.vec_sensor_tensor = st["sensors"]
    env.hand_dof_lower_limits = st["dof_lower"]
    env.hand_dof_upper_limits = st["dof_upper"]
    env.hand_default_dof_pos = torch.zeros(D)
    env.hand_dof_default_pos = st["dof_default"]
    env.hand_dof_default_vel = torch.zeros(D)
    env.num_bodies = env.rigid_body_states.shape[1]
    for name in ("hand_indices", "object_indices", "goal_object_indices", "goal_states", "goal_init_state", "object_init_state",
                 "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "hold_count_buf", "apply_reset_buf", "successes",
                 "consecutive_successes", "best_rotation_dist", "prev_targets", "cur_targets", "prev_prev_targets", "prev_actions",
                 "prev_actions_queue", "rb_forces", "random_force_prob", "random_cube_poses", "object_pose", "object_pose_wrt_wrist",
                 "obs_object_pose", "obs_object_pose_freq", "action_latency", "cube_pose_refresh_rate", "cube_pose_refresh_offset",
                 "affine_actions_scaling", "affine_actions_additive", "affine_cube_pose_scaling", "affine_cube_pose_additive",
                 "affine_dof_pos_scaling", "affine_dof_pos_additive", "cube_random_params", "hand_random_params", "gravity_vec"):
        setattr(env, name, st[name])
    env.adr_tensor_values = {k: st["adr_" + k] for k in dx.ADR_KEYS}
    env.actions = st["actions"]
    env.timeout_buf = st["timeout_buf"]
    env.obs_dict = {k: torch.zeros(N, w) for k, w in dx.OBS_KEYS}
    return env


def dextreme_random_queues(st: Dict[str, torch.Tensor], c, rnd: torch.Tensor, prnd: torch.Tensor):
    """Draw order of one Dextreme ``pre_physics_step`` and one ``post_physics_step`` for the tables ``HA_DEX_RND_*`` / ``HA_DEX_POST_RND_*``."""
    from . import dextreme_oracle as dx
    N, D = st["reset_buf"].shape[0], c.n_dofs
    goal_ids = st["reset_goal_buf"].nonzero(as_tuple=False).squeeze(-1)
    ids = st["apply_reset_buf"].nonzero(as_tuple=False).flatten()
    q: List[torch.Tensor] = []

    def target(rows, base):
        out = [rnd[rows, base:base + 4]]
        if c.apply_random_quat:
            out.append(rnd[rows, base + 4:base + 7])
        return out

    if len(goal_ids) > 0:
        q += target(goal_ids, dx.RND_GOAL)
    prob = st["random_force_prob"].clone()
    if len(ids) > 0:
        q.append(rnd[ids, dx.RND_RESET:dx.RND_RESET + 2 * D + 5])
        q += target(ids, dx.RND_RESET_GOAL)
        if c.apply_random_quat:
            q.append(rnd[ids, dx.RND_RESET + 2 * D + 5:dx.RND_RESET + 2 * D + 8])
        q.append(rnd[ids, dx.RND_FORCE_PROB])
        pr = torch.tensor(c.force_prob_range, dtype=torch.float32)
        prob[ids] = torch.exp((torch.log(pr[0]) - torch.log(pr[1])) * rnd[ids, dx.RND_FORCE_PROB] + torch.log(pr[1]))
    q += [rnd[:, dx.RND_ACTION_DELAY], rnd[:, dx.RND_ACTION_WHITE:dx.RND_ACTION_WHITE + D]]
    if c.force_scale > 0.0:
        q.append(rnd[:, dx.RND_FORCE_TEST])
        hit = (rnd[:, dx.RND_FORCE_TEST] < prob).nonzero(as_tuple=False)
        q.append(rnd[hit.squeeze(-1), dx.RND_FORCE_NORMAL:dx.RND_FORCE_NORMAL + 3].unsqueeze(1))
    pq = [prnd[:, dx.PRND_POSE_WHITE:dx.PRND_POSE_WHITE + 7], prnd[:, dx.PRND_CUBE:dx.PRND_CUBE + 5]]
    if c.apply_random_quat:
        pq.append(prnd[:, dx.PRND_CUBE_QUAT:dx.PRND_CUBE_QUAT + 3])
    pq += [prnd[:, dx.PRND_CUBE_MASK:dx.PRND_CUBE_MASK + 1], prnd[:, dx.PRND_DELAY], prnd[:, dx.PRND_DOF_WHITE:dx.PRND_DOF_WHITE + D]]
    return q, pq


def reference_dextreme_step(env, c, actions: torch.Tensor, rnd: torch.Tensor, prnd: torch.Tensor, st: Dict[str, torch.Tensor]) -> None:
    from . import dextreme_oracle as dx
    q_pre, q_post = dextreme_random_queues(st, c, rnd, prnd)
    progress = torch.clamp((rnd[:, dx.RND_PROGRESS] * c.max_episode_length).long(), max=c.max_episode_length - 1)
    old_randint = torch.randint
    torch.randint = lambda *a, **k: progress.clone()       # the all-env reset's torch.randint(0, max_episode_length) (:848-849)
    try:
        with scripted_randoms(q_pre) as q:
            env.pre_physics_step(actions)
            assert len(q) == 0, f"{len(q)} scripted random tensors were not consumed (pre)"
        with scripted_randoms(q_post) as q:
            env.post_physics_step()
            assert len(q) == 0, f"{len(q)} scripted random tensors were not consumed (post)"
    finally:
        torch.randint = old_randint
    env.timeout_buf = (env.progress_buf >= env.max_episode_length - 1) & (env.reset_buf != 0)      # adr_vec_task.py:290


def harvest_dextreme(env, st: Dict[str, torch.Tensor], c) -> Dict[str, torch.Tensor]:
    from . import dextreme_oracle as dx
    out = dict(st)
    for name in ("rew_buf", "reset_buf", "reset_goal_buf", "progress_buf", "hold_count_buf", "apply_reset_buf", "successes",
                 "consecutive_successes", "best_rotation_dist", "prev_targets", "cur_targets", "prev_prev_targets", "prev_actions",
                 "prev_actions_queue", "rb_forces", "random_force_prob", "object_pose", "object_pose_wrt_wrist", "obs_object_pose",
                 "obs_object_pose_freq", "timeout_buf", "actions", "dof_delta", "curr_rotation_dist", "random_cube_poses"):
        if hasattr(env, name):
            out[name] = getattr(env, name)
    out["frame"] = torch.tensor(env.frame, dtype=torch.long)
    if "episode_cumulative" in env.extras:
        out["episode_cumulative"] = torch.stack([env.extras["episode_cumulative"][k].to(torch.float32) for k in dx.STEP_TERMS], dim=0)
    out["obs"] = {k: env.obs_dict[k] for k, _ in dx.OBS_KEYS}
    out["obs_rows"] = torch.cat([env.obs_dict[k] for k, _ in dx.OBS_KEYS], dim=-1)
    out["obs_out"] = torch.clamp(out["obs_rows"], -c.clip_obs, c.clip_obs)
    return out


ADR_BUILTIN_KEYS = ["hand_damping", "hand_stiffness", "hand_joint_friction", "hand_armature", "hand_effort", "hand_lower", "hand_upper",
                    "hand_mass", "hand_friction_fingertips", "hand_restitution", "object_mass", "object_friction", "object_restitution"]


def attach_reference_adr(env, st: Dict[str, torch.Tensor], table) -> None:
    """ADRVecTask state (adr_vec_task.py:489-560) for ``recycle_envs`` / ``sample_adr_tensor`` / the Dextreme resampling."""
    from . import dextreme_oracle as dx
    params = {k: {"range": [0.5, 2.0], "range_path": "x", "delta": 0.01} for k in ADR_BUILTIN_KEYS}
    for k in dx.ADR_TENSOR_KEYS:
        params[k] = {"range": list(table.ranges[k]), "next_limits": list(table.next_limits[k]), "delta": 0.01}
    env.adr_params = params
    env.adr_params_keys = list(params.keys())
    env.num_adr_params = len(params)
    assert env.num_adr_params == dx.ADR_N_PARAMS
    env.adr_extended_boundary_sample = table.extended_boundary_sample
    env.worker_adr_boundary_fraction = table.boundary_fraction
    env.worker_types, env.adr_modes = st["worker_types"], st["adr_modes"]
    env.adr_tensor_values = {k: st["adr_" + k] for k in dx.ADR_TENSOR_KEYS}
    env.__dict__.pop("apply_randomizations", None)           # the class method (with the patched super) is what runs


def reference_adr_resample(env, c, table, mask: torch.Tensor, arnd: torch.Tensor, st: Dict[str, torch.Tensor]) -> None:
    """Runs the reference's ``recycle_envs`` + ``sample_adr_tensor`` (as the tail of ``adr_update`` does) and then
    ``AllegroHandDextremeADR.apply_randomizations``' own resampling (:1396-1416); only the Isaac Gym half of
    ``ADRVecTask.apply_randomizations`` is replaced."""
    from . import dextreme_oracle as dx
    mods = load_reference()
    ids = mask.nonzero(as_tuple=False).squeeze(-1)
    P = dx.ADR_N_PARAMS
    types = torch.where(arnd[ids, dx.ARND_WORKER] < table.boundary_fraction, 0, 1)
    modes = torch.clamp((arnd[ids, dx.ARND_MODE] * (2 * P)).long(), max=2 * P - 1)
    q: List[torch.Tensor] = [arnd[ids, dx.ARND_WORKER]]
    for t, _ in enumerate(dx.ADR_TENSOR_KEYS):
        n = dx.ADR_N_BUILTIN + t
        boundary = (types == 1) & ((modes == 2 * n) | (modes == 2 * n + 1))
        q.append(arnd[ids[~boundary], dx.ARND_PARAM + t])
    q += [arnd[ids, dx.ARND_LATENCY], arnd[ids, dx.ARND_RATE], arnd[ids, dx.ARND_OFFSET]]
    o = dx.ARND_AFFINE
    for dim in (c.n_dofs, c.n_dofs, 7, 7, c.n_dofs, c.n_dofs):
        q.append(arnd[ids, o:o + dim])
        o += dim

    def tensor_half(self, dr_params, randomize_buf, adr_objective=None, randomisation_callback=None):
        rows = torch.nonzero(randomize_buf, as_tuple=False).squeeze(-1)
        self.recycle_envs(rows)
        for k in self.adr_tensor_values:
            self.sample_adr_tensor(k, rows)

    cls = mods["adrv"].ADRVecTask
    saved, old_randint = cls.apply_randomizations, torch.randint
    cls.apply_randomizations = tensor_half
    torch.randint = lambda *a, **k: modes.clone()
    try:
        with scripted_randoms(q) as rest:
            type(env).apply_randomizations(env, {}, mask)
            assert len(rest) == 0, f"{len(rest)} scripted random tensors were not consumed (adr)"
    finally:
        cls.apply_randomizations, torch.randint = saved, old_randint


def load_rlgames_utils() -> types.ModuleType:
    """``isaacgymenvs/utils/rlgames_utils.py`` imported from where it lies (the real ``RLGPUAlgoObserver`` /
    ``ComplexObsRLGPUEnv``).  rl_games is not installed here: ``rl_games.common`` is stubbed with an empty
    ``AlgoObserver`` base class, which is all the observer inherits (its hooks are plain methods)."""
    if "rlu" in _loaded:
        return _loaded["rlu"]
    load_reference()

    class AlgoObserver:                                   # rl_games.common.algo_observer.AlgoObserver: hooks only
        def __init__(self):
            pass

    rg = _stub("rl_games")
    rg.common = _stub("rl_games.common", env_configurations=_Any(), vecenv=_Any())
    rg.common.algo_observer = _stub("rl_games.common.algo_observer", AlgoObserver=AlgoObserver)
    sys.modules["rl_games.common.vecenv"] = _stub("rl_games.common.vecenv", IVecEnv=object)
    sys.modules["rl_games.common.env_configurations"] = _stub("rl_games.common.env_configurations")
    rg.common.vecenv = sys.modules["rl_games.common.vecenv"]
    rg.common.env_configurations = sys.modules["rl_games.common.env_configurations"]
    sys.modules["isaacgymenvs.tasks"].isaacgym_task_map = {}
    if not hasattr(sys.modules["gym"].spaces, "Space"):
        sys.modules["gym"].spaces.Space = object              # only used in a type annotation
    _loaded["rlu"] = importlib.import_module("isaacgymenvs.utils.rlgames_utils")
    return _loaded["rlu"]
