// C-ABI entry points of the DeXtreme family (AllegroHandDextremeADR); see include/handarm_b200.h.
#include <cuda_runtime.h>

#include "../../include/handarm_b200.h"
#include "ha_dex_adr.cuh"
#include "ha_dex_post.cuh"
#include "ha_dex_pre.cuh"
#include "ha_host.h"

namespace {

int validate_dex(const HaDexParams* p, const HaDexBuffers* b) {
  HA_REQUIRE(p != nullptr && b != nullptr, "null params / buffers");
  HA_REQUIRE(p->abi_version == HA_ABI_VERSION, "HaDexParams.abi_version %d != %d", p->abi_version, HA_ABI_VERSION);
  HA_REQUIRE(p->n_envs > 0, "n_envs must be positive (got %d)", p->n_envs);
  HA_REQUIRE(p->dofs_per_env >= HA_DEX_DOFS, "dofs_per_env %d < %d", p->dofs_per_env, HA_DEX_DOFS);
  HA_REQUIRE(p->queue_len >= 1 && p->queue_len <= HA_DEX_MAX_QUEUE, "queue_len %d out of range [1, %d]", p->queue_len, HA_DEX_MAX_QUEUE);
  HA_REQUIRE(p->n_bodies > 0 && p->object_body >= 0 && p->object_body < p->n_bodies, "object_body out of range");
  HA_REQUIRE(p->palm_body >= 0 && p->palm_body < p->n_bodies, "palm_body out of range");
  for (int t = 0; t < HA_DEX_TIPS; ++t) HA_REQUIRE(p->tip_body[t] >= 0 && p->tip_body[t] < p->n_bodies, "tip_body[%d] out of range", t);
  HA_REQUIRE(p->max_episode_length >= 1, "max_episode_length must be >= 1");
  const void* need[] = {b->root_state, b->rigid_body_state, b->dof_state, b->dof_force, b->force_sensors, b->object_actor, b->goal_actor,
                        b->hand_actor, b->object_init_state, b->goal_init_state, b->action_latency, b->cube_pose_refresh_rate,
                        b->cube_pose_refresh_offset, b->adr_action_delay_prob, b->adr_affine_action_white, b->adr_affine_cube_pose_white,
                        b->adr_cube_obs_delay_prob, b->adr_affine_dof_pos_white, b->affine_actions_scaling, b->affine_actions_additive,
                        b->affine_cube_pose_scaling, b->affine_cube_pose_additive, b->affine_dof_pos_scaling, b->affine_dof_pos_additive,
                        b->cube_random_params, b->hand_random_params, b->gravity_vec, b->goal_states, b->obs_rows, b->rew_buf, b->reset_buf,
                        b->reset_goal_buf, b->progress_buf, b->hold_count_buf, b->apply_reset_buf, b->timeout_buf, b->successes,
                        b->consecutive_successes, b->best_rotation_dist, b->prev_targets, b->cur_targets, b->prev_prev_targets, b->actions,
                        b->prev_actions, b->action_queue, b->dof_delta, b->rb_forces, b->random_force_prob, b->object_pose,
                        b->obs_object_pose, b->obs_object_pose_freq, b->episode_cumulative, b->root_reset_indices, b->hand_reset_indices,
                        b->reset_counts, b->stats, b->stat_accum, b->tickets, b->dynamic};
  for (size_t i = 0; i < sizeof(need) / sizeof(need[0]); ++i) HA_REQUIRE(need[i] != nullptr, "HaDexBuffers field #%zu is null", i);
  HA_REQUIRE_ALIGNED16(b->dof_state);
  HA_REQUIRE_ALIGNED16(b->dof_force);
  HA_REQUIRE_ALIGNED16(b->force_sensors);
  HA_REQUIRE_ALIGNED16(b->obs_rows);
  HA_REQUIRE_ALIGNED16(b->actions);
  HA_REQUIRE_ALIGNED16(b->cur_targets);
  HA_REQUIRE_ALIGNED16(b->prev_targets);
  HA_REQUIRE_ALIGNED16(b->affine_actions_scaling);
  HA_REQUIRE_ALIGNED16(b->affine_actions_additive);
  HA_REQUIRE_ALIGNED16(b->affine_cube_pose_scaling);
  HA_REQUIRE_ALIGNED16(b->affine_cube_pose_additive);
  HA_REQUIRE_ALIGNED16(b->affine_dof_pos_scaling);
  HA_REQUIRE_ALIGNED16(b->affine_dof_pos_additive);
  HA_REQUIRE_ALIGNED16(b->cube_random_params);
  HA_REQUIRE_ALIGNED16(b->hand_random_params);
  HA_REQUIRE_ALIGNED16(b->gravity_vec);
  HA_REQUIRE_ALIGNED16(b->prev_actions);
  HA_REQUIRE_ALIGNED16(b->action_queue);
  HA_REQUIRE_ALIGNED16(b->dof_delta);
  HA_REQUIRE_ALIGNED16(b->obs_object_pose);
  HA_REQUIRE_ALIGNED16(b->obs_object_pose_freq);
  if (p->root_regular)
    HA_REQUIRE(p->actors_per_env > 0 && p->object_slot >= 0 && p->object_slot < p->actors_per_env && p->goal_slot >= 0 &&
                   p->goal_slot < p->actors_per_env, "root_regular needs 0 <= object_slot, goal_slot < actors_per_env");
  for (int v = 0; v < 2; ++v) {
    HA_REQUIRE(p->n_view_cols[v] >= 0 && p->n_view_cols[v] <= HA_DEX_OBS_DIM, "n_view_cols[%d] = %d out of range", v, p->n_view_cols[v]);
    HA_REQUIRE(p->n_view_cols[v] == 0 || (b->view_cols[v] != nullptr && b->view_rows[v] != nullptr), "view %d: view_cols / view_rows is null", v);
  }
  return HA_OK;
}

int launch_dex_post(const HaDexParams* p, const HaDexBuffers* b, const float* ext_rand, int mode, cudaStream_t st) {
  static ha::PerDeviceFlag flag;
  bool& configured = flag.here();
  const size_t smem = ha::dex_post_smem_bytes();
  if (!configured) {
    int rc = ha::check_cuda(cudaFuncSetAttribute(ha::dex_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                            "cudaFuncSetAttribute(dex_post_kernel)");
    if (rc) return rc;
    configured = true;
  }
  const int grid = (p->n_envs + ha::kDexTile - 1) / ha::kDexTile;
  return ha::check_cuda(ha::launch(ha::dex_post_kernel, grid, ha::kDexThreads, smem, st, *p, *b, ext_rand, mode), "dex_post_kernel launch");
}

}  // namespace

extern "C" {

size_t ha_sizeof_dex_params(void) { return sizeof(HaDexParams); }
size_t ha_sizeof_dex_buffers(void) { return sizeof(HaDexBuffers); }
size_t ha_sizeof_dex_adr_table(void) { return sizeof(HaDexAdrTable); }
size_t ha_sizeof_dex_adr_buffers(void) { return sizeof(HaDexAdrBuffers); }

int ha_dex_adr_resample(const HaDexParams* p, const HaDexBuffers* b, const HaDexAdrTable* t, const HaDexAdrBuffers* ab, const float* ext_rand,
                        ha_stream_t stream) {
  int rc = validate_dex(p, b);
  if (rc) return rc;
  HA_REQUIRE(t != nullptr && ab != nullptr, "null ADR table / buffers");
  HA_REQUIRE(t->n_tensor >= 0 && t->n_tensor <= HA_DEX_ADR_MAX_PARAMS, "n_tensor %d out of range", t->n_tensor);
  HA_REQUIRE(t->n_params_total >= t->n_tensor && t->n_params_total > 0, "n_params_total %d < n_tensor %d", t->n_params_total, t->n_tensor);
  for (int k = 0; k < t->n_tensor; ++k) {
    HA_REQUIRE(ab->values[k] != nullptr, "HaDexAdrBuffers.values[%d] is null", k);
    HA_REQUIRE(t->param_index[k] >= 0 && t->param_index[k] < t->n_params_total, "param_index[%d] out of range", k);
  }
  HA_REQUIRE(ab->worker_types != nullptr && ab->adr_modes != nullptr, "worker_types / adr_modes is null");
  HA_REQUIRE(t->idx_action_latency >= 0 && t->idx_action_latency < t->n_tensor && t->idx_cube_pose_refresh_rate >= 0 &&
                 t->idx_cube_pose_refresh_rate < t->n_tensor, "action_latency / cube_pose_refresh_rate positions out of range");
  for (int a = 0; a < 6; ++a) HA_REQUIRE(t->idx_affine[a] >= 0 && t->idx_affine[a] < t->n_tensor, "idx_affine[%d] out of range", a);
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  const int grid = (p->n_envs + ha::kDexThreads - 1) / ha::kDexThreads;
  return ha::check_cuda(ha::launch(ha::dex_adr_kernel, grid, ha::kDexThreads, 0, (cudaStream_t)stream, *p, *b, *t, *ab, ext_rand), "dex_adr_kernel launch");
}

void ha_dex_fill_derived(HaDexParams* p) {
  if (!p) return;
  for (int d = 0; d < HA_DEX_DOFS; ++d) {
    const float range = p->dof_upper[d] - p->dof_lower[d];
    p->dof_range_rcp[d] = range != 0.0f ? (float)(1.0 / (double)range) : 0.0f;
  }
}

// SURVEY.md 8d counting rule: every needed input element once, every output element once, int64 surface buffers at 8 B
size_t ha_dex_post_bytes_per_env_step(const HaDexParams* p) {
  if (!p) return 0;
  // reads: DOF state 128, DOF forces 64, force sensors 96, object row 52, goal pose 28, palm pose 28, 4 fingertip rows 208, actions 64,
  //        cur + prev targets 128, six affine tensors 312, cube / hand / gravity parameters 28, object force row 12,
  //        latency + refresh rate + offset 24, four ADR scalars 16, reset / reset_goal / progress / hold 32, successes + best 8, two latches 56
  const size_t rd = 128 + 64 + 96 + 52 + 28 + 28 + 208 + 64 + 128 + 312 + 28 + 12 + 24 + 16 + 32 + 8 + 56;
  // writes: 285-float row 1140, reward 4, four int64 32, successes + best 8, prev_targets 64, object pose 28, two latches 56, 8 reward terms 32
  const size_t wr = 4 * HA_DEX_OBS_DIM + 4 + 32 + 8 + 64 + 28 + 56 + 4 * HA_DEX_N_STEP_TERMS;
  return rd + wr;
}

size_t ha_dex_pre_bytes_per_env_step(const HaDexParams* p) {
  if (!p) return 0;
  // actions r 64, queue slot w 64, delayed slot r 64, prev_actions r + w 128, action affine 128, prev_targets r 64, cur_targets w 64,
  // actions copy w 64, dof_delta w 64, three int64 masks + progress 32, latency 8, two ADR scalars 8, force row r + w 24, force prob 4
  return 64 + 64 + 64 + 128 + 128 + 64 + 64 + 64 + 64 + 32 + 8 + 8 + (p->force_enabled ? 28 : 0);
}

size_t ha_dex_bytes_per_env_step(const HaDexParams* p) { return ha_dex_post_bytes_per_env_step(p) + ha_dex_pre_bytes_per_env_step(p); }

int ha_dex_pre_step(const HaDexParams* p, const HaDexBuffers* b, const float* actions, const float* rna_perturbation, const float* ext_rand,
                    ha_stream_t stream) {
  int rc = validate_dex(p, b);
  if (rc) return rc;
  HA_REQUIRE(actions != nullptr, "actions is null");
  HA_REQUIRE_ALIGNED16(actions);
  if (rna_perturbation) HA_REQUIRE_ALIGNED16(rna_perturbation);
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  HA_REQUIRE(rna_perturbation == nullptr || b->adr_rna_alpha != nullptr, "adr_rna_alpha is required with rna_perturbation");
  const int grid = (p->n_envs + ha::kDexTile - 1) / ha::kDexTile;
  return ha::check_cuda(ha::launch(ha::dex_pre_kernel, grid, ha::kDexThreads, 0, (cudaStream_t)stream, *p, *b, actions, rna_perturbation, ext_rand),
                        "dex_pre_kernel launch");
}

int ha_dex_post_step(const HaDexParams* p, const HaDexBuffers* b, const float* ext_rand_post, ha_stream_t stream) {
  int rc = validate_dex(p, b);
  if (rc) return rc;
  HA_REQUIRE(p->use_fast_rng || ext_rand_post != nullptr, "ext_rand_post is required unless use_fast_rng is set");
  return launch_dex_post(p, b, ext_rand_post, ha::kDexPostFull, (cudaStream_t)stream);
}

int ha_dex_observe(const HaDexParams* p, const HaDexBuffers* b, const float* ext_rand_post, ha_stream_t stream) {
  int rc = validate_dex(p, b);
  if (rc) return rc;
  HA_REQUIRE(p->use_fast_rng || ext_rand_post != nullptr, "ext_rand_post is required unless use_fast_rng is set");
  return launch_dex_post(p, b, ext_rand_post, ha::kDexPostObserveOnly, (cudaStream_t)stream);
}

}  // extern "C"
