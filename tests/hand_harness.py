"""Test harness: move hand oracle state dicts in and out of the CUDA ShadowHand / AllegroHand task (tests only)."""
from __future__ import annotations

from typing import Dict

import torch

import handarm_b200.config as hcfg
from handarm_b200.sim import ALLEGRO_HAND_LAYOUT, SHADOW_HAND_LAYOUT, SyntheticSim
from handarm_b200.tasks.hand import AllegroHand, ShadowHand

DIRECT = ["goal_states", "goal_init_state", "object_init_state", "obs_buf", "rew_buf", "reset_buf", "reset_goal_buf", "progress_buf",
          "randomize_buf", "timeout_buf", "successes", "consecutive_successes", "prev_targets", "cur_targets", "actions", "rb_forces",
          "random_force_prob"]


def overrides_from_consts(c) -> Dict[str, object]:
    return {"env.observationType": c.obs_type, "env.asymmetric_observations": c.asymmetric, "env.episodeLength": c.max_episode_length,
            "env.distRewardScale": c.dist_reward_scale, "env.rotRewardScale": c.rot_reward_scale, "env.rotEps": c.rot_eps,
            "env.actionPenaltyScale": c.action_penalty_scale, "env.successTolerance": c.success_tolerance,
            "env.reachGoalBonus": c.reach_goal_bonus, "env.fallDistance": c.fall_dist, "env.fallPenalty": c.fall_penalty,
            "env.maxConsecutiveSuccesses": c.max_consecutive_successes, "env.averFactor": c.av_factor,
            "env.useRelativeControl": c.use_relative_control, "env.dofSpeedScale": c.dof_speed_scale,
            "env.actionsMovingAverage": c.act_moving_average, "env.forceScale": c.force_scale, "env.forceDecay": c.force_decay,
            "env.forceDecayInterval": c.force_decay_interval, "env.forceProbRange": list(c.force_prob_range),
            "env.resetPositionNoise": c.reset_position_noise, "env.resetDofPosRandomInterval": c.reset_dof_pos_noise,
            "env.resetDofVelRandomInterval": c.reset_dof_vel_noise, "env.clipObservations": c.clip_obs, "sim.dt": c.dt,
            "env.objectType": "pen" if c.ignore_z_rot else "block"}


def make_task(n: int, c, device: str = "cuda:0"):
    name, cls, layout = ("ShadowHand", ShadowHand, SHADOW_HAND_LAYOUT) if c.kind == "shadow" else ("AllegroHand", AllegroHand,
                                                                                                  ALLEGRO_HAND_LAYOUT)
    cfg = hcfg.load_task_config(name, num_envs=n, sim_device=device, overrides=overrides_from_consts(c))
    sim = SyntheticSim(n, layout, device, seed=1, snapshots=1)
    task = cls(cfg, device, device, 0, True, sim=sim)
    assert task.num_obs == c.obs_dim and task.num_states == c.states_dim
    return task


def load_state(task, st: Dict[str, torch.Tensor]) -> None:
    dev = task.device
    for k in DIRECT:
        getattr(task, k).copy_(st[k].to(dev))
    if task.num_states > 0:
        task.states_buf.copy_(st["states_buf"].to(dev))
    task.object_rb_masses.fill_(0.1)
    task.sim.root_rows().copy_(st["root_state"].view(task.num_envs, -1, 13).to(dev))
    task.sim.body_rows().copy_(st["rigid_body_state"].view(task.num_envs, -1, 13).to(dev))
    task.sim.dof_state.copy_(st["dof_state"].view(-1, 2).to(dev))
    task.sim.dof_force.copy_(st["dof_force"].to(dev))
    if task.is_shadow:
        task.sim.force_sensors.copy_(st["sensors"].to(dev))


def dump_state(task) -> Dict[str, torch.Tensor]:
    torch.cuda.synchronize()
    out = {k: getattr(task, k).detach().cpu().clone() for k in DIRECT}
    out["states_buf"] = task.states_buf.cpu().clone()
    out["root_state"] = task.sim.root_rows().cpu().reshape(-1, 13).clone()
    out["dof_state"] = task.sim.dof_state.view(task.num_envs, -1, 2).cpu().clone()
    n_root, n_hand = int(task.reset_counts[0]), int(task.reset_counts[1])
    out["root_rows"] = task.root_reset_indices[:n_root].cpu().clone()
    out["hand_rows"] = task.hand_reset_indices[:n_hand].cpu().clone()
    out["stats"] = task.stats.cpu().clone()
    return out
