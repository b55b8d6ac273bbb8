"""Test harness: move oracle state dicts in and out of the CUDA task (tests only)."""
from __future__ import annotations

from typing import Dict

import torch

import handarm_b200.config as hcfg
from handarm_b200.sim import KUKA_LAYOUT, SyntheticSim
from handarm_b200.tasks.allegro_kuka import AllegroKukaRegrasping, AllegroKukaReorientation, AllegroKukaThrow

SUBTASK_CLASSES = {"reorientation": AllegroKukaReorientation, "regrasping": AllegroKukaRegrasping, "throw": AllegroKukaThrow}

# state-dict key -> task attribute holding the same tensor
DIRECT = ["goal_states", "object_init_state", "object_keypoint_offsets", "object_scales", "finger_rew_coeffs", "obs_buf", "rew_buf",
          "reset_buf", "reset_goal_buf", "progress_buf", "randomize_buf", "timeout_buf", "successes", "prev_episode_successes",
          "true_objective", "prev_episode_true_objective", "near_goal_steps", "lifted_object", "closest_keypoint_max_dist",
          "closest_fingertip_dist", "furthest_hand_dist", "prev_targets", "cur_targets", "actions", "rb_forces", "random_force_prob"]


STATE_LAYOUT = "isaacgym"      # tests switch this to "soa" to run the same cases on the component-major producer layout


def make_task(n: int, device: str = "cuda:0", overrides=None, snapshots: int = 1, state_layout=None,
              subtask: str = "reorientation") -> AllegroKukaReorientation:
    cfg = hcfg.load_task_config("AllegroKuka", num_envs=n, sim_device=device, overrides=overrides,
                                subtask=None if subtask == "reorientation" else subtask)
    sim = SyntheticSim(n, KUKA_LAYOUT, device, seed=1, snapshots=snapshots, state_layout=state_layout or STATE_LAYOUT)
    return SUBTASK_CLASSES[subtask](cfg, device, device, 0, True, sim=sim)


def load_state(task, st: Dict[str, torch.Tensor]) -> None:
    dev = task.device
    for k in DIRECT:
        getattr(task, k).copy_(st[k].to(dev))
    task._rewards_episode.copy_(st["rewards_episode"].to(dev))
    task._episode_cumulative.copy_(st["episode_cumulative"].to(dev))
    task._object_init_z.copy_(st["object_init_state"][:, 2].to(dev))
    task.sim.root_rows().copy_(st["root_state"].view(task.num_envs, -1, 13).to(dev))
    task.sim.body_rows().copy_(st["rigid_body_state"].view(task.num_envs, -1, 13).to(dev))
    task.sim.dof_state.copy_(st["dof_state"].view(-1, 2).to(dev))


def dump_state(task) -> Dict[str, torch.Tensor]:
    torch.cuda.synchronize()
    out = {k: getattr(task, k).detach().cpu().clone() for k in DIRECT}
    out["rewards_episode"] = task._rewards_episode.cpu().clone()
    out["episode_cumulative"] = task._episode_cumulative.cpu().clone()
    out["root_state"] = task.sim.root_rows().cpu().reshape(-1, 13).clone()
    out["rigid_body_state"] = task.sim.body_rows().cpu().reshape(-1, 13).clone()
    out["dof_state"] = task.sim.dof_state.view(task.num_envs, -1, 2).cpu().clone()
    n_root, n_hand = int(task.reset_counts[0]), int(task.reset_counts[1])
    out["root_rows"] = task.root_reset_indices[:n_root].cpu().clone()
    out["hand_rows"] = task.hand_reset_indices[:n_hand].cpu().clone()
    out["stats"] = task.stats.cpu().clone()
    return out
