"""Synthetic stand-in for the closed Isaac Gym / PhysX binary.

The north star treats the PhysX step as an opaque producer of the actor-root, rigid-body and DOF
state tensors.  ``SyntheticSim`` allocates those tensors with Isaac Gym's shapes and strides
(``allegro_kuka_base.py:280-295``: root ``[N*A, 13]`` = pos3 quat4(xyzw) linvel3 angvel3, rigid-body
``[N*B, 13]``, DOF ``[N*D, 2]`` = pos, vel), fills ``S`` rotating snapshots from a seeded generator
(distributions of SURVEY.md 8d) and records the calls a task makes into the simulator
(``set_*_indexed``, ``apply_rigid_body_force_tensors`` ...) instead of executing them.
"""
from __future__ import annotations

import dataclasses
import math
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

# Joint limits of assets/urdf/kuka_allegro_description/kuka_allegro_touch_sensor.urdf (7 iiwa + 16 allegro)
KUKA_ALLEGRO_LOWER = [-2.96705972839, -2.09439510239, -2.96705972839, -2.09439510239, -2.96705972839, -2.09439510239,
                      -3.05432619099] + [-0.558488888889, -0.279244444444, -0.279244444444, -0.279244444444] * 3 + [
                          0.279244444444, -0.331602777778, -0.279244444444, -0.279244444444]
KUKA_ALLEGRO_UPPER = [2.96705972839, 2.09439510239, 2.96705972839, 2.09439510239, 2.96705972839, 2.09439510239,
                      3.05432619099] + [0.558488888889, 1.727825, 1.727825, 1.727825] * 3 + [1.57075, 1.15188333333, 1.727825,
                                                                                             1.76273055556]


@dataclasses.dataclass
class ActorLayout:
    """Index tables Isaac Gym would assign (parameters of the harness, not constants)."""
    actors_per_env: int
    bodies_per_env: int
    dofs_per_env: int
    hand_actor: Sequence[int]        # actor slot(s) of the arm(s) inside an env
    object_actor: int
    goal_actor: int                  # -1 if the task has no goal actor
    palm_body: Sequence[int]
    tip_body: Sequence[int]
    object_body: int


# collapse_fixed_joints=True body numbering, SURVEY.md 8 table / allegro_kuka_base.py:649-656
KUKA_LAYOUT = ActorLayout(actors_per_env=4, bodies_per_env=27, dofs_per_env=23, hand_actor=(0,), object_actor=1, goal_actor=3,
                          palm_body=(7,), tip_body=(11, 15, 19, 23), object_body=24)


# two arms (allegro_kuka_two_arms.py:652-680): arm0 (24 bodies), arm1 (24), object, table, goal ball
KUKA2_LAYOUT = ActorLayout(actors_per_env=5, bodies_per_env=51, dofs_per_env=46, hand_actor=(0, 1), object_actor=2, goal_actor=4,
                           palm_body=(7, 31), tip_body=(11, 15, 19, 23, 35, 39, 43, 47), object_body=48)

# ShadowHand (mjcf/open_ai_assets/hand/shadow_hand.xml): 26 hand bodies + object + goal, 24 joints, fingertips = the five
# *distal bodies; AllegroHand (allegro_touch_sensor.urdf): 17 hand bodies + object + goal, 16 joints.  Actors: hand, object, goal.
SHADOW_HAND_LAYOUT = ActorLayout(actors_per_env=3, bodies_per_env=28, dofs_per_env=24, hand_actor=(0,), object_actor=1, goal_actor=2,
                                 palm_body=(), tip_body=(7, 11, 15, 20, 25), object_body=26)
ALLEGRO_HAND_LAYOUT = ActorLayout(actors_per_env=3, bodies_per_env=19, dofs_per_env=16, hand_actor=(0,), object_actor=1, goal_actor=2,
                                  palm_body=(), tip_body=(), object_body=17)
# joint limits of assets/mjcf/open_ai_assets/hand/robot.xml (WRJ1 WRJ0 FFJ3..0 MFJ3..0 RFJ3..0 LFJ4..0 THJ4..0)
SHADOW_HAND_LOWER = [-0.489, -0.698, -0.349, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0, 0.0, -0.349, 0.0, 0.0, 0.0,
                     -1.047, 0.0, -0.209, -0.524, -1.571]
SHADOW_HAND_UPPER = [0.14, 0.489, 0.349, 1.571, 1.571, 1.571, 0.349, 1.571, 1.571, 1.571, 0.349, 1.571, 1.571, 1.571, 0.785, 0.349,
                     1.571, 1.571, 1.571, 1.047, 1.222, 0.209, 0.524, 0.0]
ALLEGRO_HAND_LOWER = KUKA_ALLEGRO_LOWER[7:]
ALLEGRO_HAND_UPPER = KUKA_ALLEGRO_UPPER[7:]
# actuated joints of shared.xml (all but FFJ0 / MFJ0 / RFJ0 / LFJ0, which are tendon-coupled)
SHADOW_HAND_ACTUATED = [0, 1, 2, 3, 4, 6, 7, 8, 10, 11, 12, 14, 15, 16, 17, 19, 20, 21, 22, 23]


def cuboid_scales(small: bool = True, big: bool = True, sticks: bool = True) -> List[List[float]]:
    """Object scale table the reference derives from its generated cuboid assets.

    Restates the enumeration of ``tasks/allegro_kuka/generate_cuboids.py:36-140`` (volume windows
    and filters) without the URDF templating, then applies the fixed-seed shuffle of
    ``allegro_kuka_base.py:416-428`` so env ``i`` gets the same ``object_scales[i % len]``.
    """
    def enum(scales, vmin, vmax, filters, start):
        out, idx = [], start
        for x in scales:
            for y in scales:
                for z in scales:
                    vol = x * y * z / 1e6
                    if vol > vmax or vol < vmin:
                        continue
                    srt = sorted([x, y, z])
                    if any(f(srt) for f in filters):
                        continue
                    out.append((f"{idx:03d}_cube_{x}_{y}_{z}", [x / 100, y / 100, z / 100]))
                    idx += 1
        return out

    thin = lambda s: s[0] * 3 <= s[1]
    non_elongated = lambda s: s[2] <= s[0] * 3 or s[2] <= s[1] * 3
    # every generator restarts its file index at 0 and files with equal names overwrite each other
    files: Dict[str, List[float]] = {}
    files.update(enum([100], 1.0, 1.0, [], 0))
    if small:
        files.update(enum([100, 50, 66, 75, 90, 110, 125, 150, 175, 200, 250, 300], 1.0, 2.5, [], 0))
    if big:
        files.update(enum([100, 125, 150, 200, 250, 300, 350], 2.5, 15.0, [thin], 0))
    if sticks:
        files.update(enum([100, 50, 75, 200, 300, 400, 500, 600], 2.5, 6.0, [thin, non_elongated], 0))
    items = [(k, files[k]) for k in sorted(files)]
    rng = np.random.default_rng(42)
    rng.shuffle(items)
    return [s for _, s in items]


def _rand_unit_quat(gen: torch.Generator, shape) -> torch.Tensor:
    q = torch.randn(*shape, 4, generator=gen)
    return q / q.norm(dim=-1, keepdim=True).clamp_min(1e-6)


class SyntheticSim:
    """Owns the simulator-side tensors and a ring of ``S`` pre-generated snapshots."""

    def __init__(self, num_envs: int, layout: ActorLayout, device: str = "cuda:0", seed: int = 1234, snapshots: int = 1,
                 state_layout: str = "isaacgym"):
        """``state_layout``: ``"isaacgym"`` = the tensor API's AoS rows (root ``[N*A, 13]``, rigid-body ``[N*B, 13]``);
        ``"soa"`` = component-major planes (root ``[A, 13, N]``, rigid-body ``[B, 13, N]``, ``HA_STATE_SOA``) for producers
        that can emit them.  ``dof_state`` is ``[N*D, 2]`` in both."""
        if state_layout not in ("isaacgym", "soa"):
            raise ValueError(f"unknown state_layout {state_layout!r}")
        self.state_layout = state_layout
        self.num_envs = int(num_envs)
        self.layout = layout
        self.device = torch.device(device)
        self.seed = seed
        self.n_snapshots = int(snapshots)
        self.cursor = 0
        self.calls: List[Tuple[str, tuple]] = []
        self.record_calls = True
        N, L = self.num_envs, layout
        if state_layout == "soa":
            self.root_snap = [torch.zeros(L.actors_per_env, 13, N, device=self.device) for _ in range(self.n_snapshots)]
            self.rb_snap = [torch.zeros(L.bodies_per_env, 13, N, device=self.device) for _ in range(self.n_snapshots)]
        else:
            self.root_snap = [torch.zeros(N * L.actors_per_env, 13, device=self.device) for _ in range(self.n_snapshots)]
            self.rb_snap = [torch.zeros(N * L.bodies_per_env, 13, device=self.device) for _ in range(self.n_snapshots)]
        self.dof_snap = [torch.zeros(N * L.dofs_per_env, 2, device=self.device) for _ in range(self.n_snapshots)]
        # acquire_dof_force_tensor / acquire_force_sensor_tensor (hand tasks, full_state observations)
        self.dof_force_snap = [torch.zeros(N, L.dofs_per_env, device=self.device) for _ in range(self.n_snapshots)]
        self.sensor_snap = [torch.zeros(N, 6 * max(len(L.tip_body), 1), device=self.device) for _ in range(self.n_snapshots)]

    # ---- the acquire_* surface -----------------------------------------------------------
    @property
    def root_state(self) -> torch.Tensor:
        return self.root_snap[self.cursor]

    @property
    def rigid_body_state(self) -> torch.Tensor:
        return self.rb_snap[self.cursor]

    @property
    def dof_state(self) -> torch.Tensor:
        return self.dof_snap[self.cursor]

    @property
    def dof_force(self) -> torch.Tensor:
        return self.dof_force_snap[self.cursor]

    @property
    def force_sensors(self) -> torch.Tensor:
        return self.sensor_snap[self.cursor]

    # layout-independent [N, rows, 13] views (strided in the "soa" layout; writes go through)
    def root_rows(self, snapshot: Optional[int] = None) -> torch.Tensor:
        t = self.root_snap[self.cursor if snapshot is None else snapshot]
        return t.permute(2, 0, 1) if self.state_layout == "soa" else t.view(self.num_envs, -1, 13)

    def body_rows(self, snapshot: Optional[int] = None) -> torch.Tensor:
        t = self.rb_snap[self.cursor if snapshot is None else snapshot]
        return t.permute(2, 0, 1) if self.state_layout == "soa" else t.view(self.num_envs, -1, 13)

    def simulate(self) -> None:
        """One opaque physics step: expose the next pre-generated snapshot (pointer swap, no copy)."""
        self.cursor = (self.cursor + 1) % self.n_snapshots

    def record(self, name: str, *args: Any) -> None:
        if self.record_calls:
            self.calls.append((name, args))

    # ---- synthetic content ------------------------------------------------------------
    def fill_kuka(self, object_init_z: float = 0.63, goal_states: Optional[torch.Tensor] = None,
                  lower: Sequence[float] = KUKA_ALLEGRO_LOWER, upper: Sequence[float] = KUKA_ALLEGRO_UPPER) -> None:
        """Seeded AllegroKuka-like state for every snapshot (SURVEY.md 8d distributions)."""
        N, L = self.num_envs, self.layout
        lo = torch.tensor(lower, dtype=torch.float32)
        hi = torch.tensor(upper, dtype=torch.float32)
        D = L.dofs_per_env
        reps = D // lo.numel()
        lo, hi = lo.repeat(reps), hi.repeat(reps)
        for s in range(self.n_snapshots):
            gen = torch.Generator().manual_seed(self.seed * 1000003 + s)
            root = torch.zeros(N, L.actors_per_env, 13)
            root[..., 0:3] = torch.tensor([0.0, 0.0, object_init_z]) + 0.1 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            root[..., 3:7] = _rand_unit_quat(gen, (N, L.actors_per_env))
            root[..., 7:10] = 0.5 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            root[..., 10:13] = 2.0 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            # ~1% of the objects have fallen below the table (z < 0.1), ~50% are lifted
            fallen = torch.rand(N, generator=gen) < 0.01
            root[fallen, L.object_actor, 2] = 0.05
            rb = torch.zeros(N, L.bodies_per_env, 13)
            # bodies live around the hand workspace; fingertips within ~0.1-0.3 m of the object
            rb[..., 0:3] = root[:, L.object_actor:L.object_actor + 1, 0:3] + 0.12 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            rb[..., 3:7] = _rand_unit_quat(gen, (N, L.bodies_per_env))
            rb[..., 7:10] = 0.5 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            rb[..., 10:13] = 2.0 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            # a few hands drift far from the object (reorientation's > 1.5 m reset rule)
            far = torch.rand(N, generator=gen) < 0.003
            rb[far, :, 0] += 2.0
            rb[:, L.object_body] = root[:, L.object_actor]
            dof = torch.zeros(N, D, 2)
            dof[..., 0] = lo + (hi - lo) * torch.rand(N, D, generator=gen)
            dof[..., 1] = torch.randn(N, D, generator=gen)
            if goal_states is not None:
                # 2% of the envs sit (almost) on the goal pose so the success branch fires
                hit = torch.rand(N, generator=gen) < 0.02
                g = goal_states.detach().cpu()
                root[hit, L.object_actor, 0:7] = g[hit, 0:7] + 1e-3 * torch.randn(int(hit.sum()), 7, generator=gen)
                rb[:, L.object_body] = root[:, L.object_actor]
            self.root_rows(s).copy_(root)
            self.body_rows(s).copy_(rb)
            self.dof_snap[s].copy_(dof.view(-1, 2))

    def fill_hand(self, object_init: Sequence[float], goal_states: Optional[torch.Tensor], lower: Sequence[float],
                  upper: Sequence[float]) -> None:
        """Seeded in-hand state for every snapshot (ShadowHand / AllegroHand): the object hovers around its start pose,
        ~3 % of the objects have fallen (goal_dist >= fallDistance), ~2 % sit at the goal orientation."""
        N, L = self.num_envs, self.layout
        lo, hi = torch.tensor(lower, dtype=torch.float32), torch.tensor(upper, dtype=torch.float32)
        D = L.dofs_per_env
        for s in range(self.n_snapshots):
            gen = torch.Generator().manual_seed(self.seed * 1000003 + s)
            root = torch.zeros(N, L.actors_per_env, 13)
            root[..., 0:3] = torch.tensor(list(object_init), dtype=torch.float32) + 0.03 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            root[..., 3:7] = _rand_unit_quat(gen, (N, L.actors_per_env))
            root[..., 7:10] = 0.5 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            root[..., 10:13] = 2.0 * torch.randn(N, L.actors_per_env, 3, generator=gen)
            fallen = torch.rand(N, generator=gen) < 0.03
            root[fallen, L.object_actor, 2] -= 0.4
            if goal_states is not None:
                hit = torch.rand(N, generator=gen) < 0.02
                g = goal_states.detach().cpu()
                q = g[hit, 3:7] + 1e-3 * torch.randn(int(hit.sum()), 4, generator=gen)
                root[hit, L.object_actor, 3:7] = q / q.norm(dim=-1, keepdim=True)
            rb = torch.zeros(N, L.bodies_per_env, 13)
            rb[..., 0:3] = root[:, L.object_actor:L.object_actor + 1, 0:3] + 0.08 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            rb[..., 3:7] = _rand_unit_quat(gen, (N, L.bodies_per_env))
            rb[..., 7:10] = 0.5 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            rb[..., 10:13] = 2.0 * torch.randn(N, L.bodies_per_env, 3, generator=gen)
            rb[:, L.object_body] = root[:, L.object_actor]
            dof = torch.zeros(N, D, 2)
            dof[..., 0] = lo + (hi - lo) * torch.rand(N, D, generator=gen)
            dof[..., 1] = 3.0 * torch.randn(N, D, generator=gen)
            self.root_rows(s).copy_(root)
            self.body_rows(s).copy_(rb)
            self.dof_snap[s].copy_(dof.view(-1, 2))
            self.dof_force_snap[s].copy_(0.3 * torch.randn(N, D, generator=gen))
            self.sensor_snap[s].copy_(0.2 * torch.randn(self.sensor_snap[s].shape, generator=gen))
