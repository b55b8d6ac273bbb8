// C-ABI entry points of the in-hand reorientation family (ShadowHand, AllegroHand); see include/handarm_b200.h.
#include <cuda_runtime.h>

#include "../../include/handarm_b200.h"
#include "ha_hand_post.cuh"
#include "ha_hand_pre.cuh"
#include "ha_host.h"

namespace {

int validate_hand(const HaHandParams* p, const HaHandBuffers* b) {
  HA_REQUIRE(p != nullptr && b != nullptr, "null params / buffers");
  HA_REQUIRE(p->abi_version == HA_ABI_VERSION, "HaHandParams.abi_version %d != %d", p->abi_version, HA_ABI_VERSION);
  HA_REQUIRE(p->n_envs > 0, "n_envs must be positive (got %d)", p->n_envs);
  HA_REQUIRE(p->n_dofs > 0 && p->n_dofs <= HA_HAND_MAX_DOFS && p->dofs_per_env >= p->n_dofs, "n_dofs %d / dofs_per_env %d out of range",
             p->n_dofs, p->dofs_per_env);
  HA_REQUIRE(p->n_actions > 0 && p->n_actions <= p->n_dofs, "n_actions %d out of range", p->n_actions);
  for (int a = 0; a < p->n_actions; ++a)
    HA_REQUIRE(p->actuated_dof[a] >= 0 && p->actuated_dof[a] < p->n_dofs, "actuated_dof[%d] out of range", a);
  HA_REQUIRE(p->n_tips >= 0 && p->n_tips <= HA_HAND_MAX_TIPS, "n_tips %d out of range", p->n_tips);
  HA_REQUIRE(!p->is_shadow || p->n_tips > 0, "ShadowHand observes fingertips: n_tips must be > 0");
  HA_REQUIRE(p->obs_type >= HA_HAND_OBS_OPENAI && p->obs_type <= HA_HAND_OBS_FULL_STATE, "unknown obs_type %d", p->obs_type);
  HA_REQUIRE(p->is_shadow || p->obs_type != HA_HAND_OBS_OPENAI, "the openai layout exists for ShadowHand only");
  const int want = ha_hand_obs_dim(p->is_shadow, p->obs_type);
  HA_REQUIRE(p->obs_dim == want && want <= HA_HAND_MAX_OBS, "obs_dim %d does not match the layout (%d)", p->obs_dim, want);
  const int full = ha_hand_obs_dim(p->is_shadow, HA_HAND_OBS_FULL_STATE);
  HA_REQUIRE(p->states_dim == 0 || p->states_dim == full, "states_dim must be 0 or %d (got %d)", full, p->states_dim);
  // the layouts hard-code the widths of the reference's slices (shadow_hand.py:498-584): check the dims that produce them
  const int T = p->is_shadow ? p->n_tips : 0;
  const int w_full = 3 * p->n_dofs + 13 + 7 + 4 + 13 * T + 6 * T + p->n_actions;
  HA_REQUIRE(w_full == full, "dims (dofs %d, tips %d, actions %d) do not give the %d-wide full_state row", p->n_dofs, T, p->n_actions, full);
  HA_REQUIRE(p->n_bodies > 0 && p->object_body >= 0 && p->object_body < p->n_bodies, "object_body out of range");
  for (int t = 0; t < T; ++t) HA_REQUIRE(p->tip_body[t] >= 0 && p->tip_body[t] < p->n_bodies, "tip_body[%d] out of range", t);
  HA_REQUIRE(p->up_axis_idx == 1 || p->up_axis_idx == 2, "up_axis_idx must be 1 or 2");
  const void* need[] = {b->root_state, b->rigid_body_state, b->dof_state, b->object_actor, b->goal_actor, b->hand_actor,
                        b->object_init_state, b->goal_init_state, b->goal_states, b->obs_buf, b->rew_buf, b->reset_buf, b->reset_goal_buf,
                        b->progress_buf, b->timeout_buf, b->successes, b->consecutive_successes, b->prev_targets, b->cur_targets,
                        b->actions, b->rb_forces, b->random_force_prob, b->root_reset_indices, b->hand_reset_indices, b->reset_counts,
                        b->stats, b->stat_accum, b->tickets, b->dynamic};
  for (size_t i = 0; i < sizeof(need) / sizeof(need[0]); ++i) HA_REQUIRE(need[i] != nullptr, "HaHandBuffers field #%zu is null", i);
  const bool need_frc = p->obs_type == HA_HAND_OBS_FULL_STATE || p->states_dim > 0;
  HA_REQUIRE(!need_frc || b->dof_force != nullptr, "dof_force is required by full_state / asymmetric observations");
  HA_REQUIRE(!(need_frc && T > 0) || b->force_sensors != nullptr, "force_sensors is required by ShadowHand full_state / asymmetric observations");
  HA_REQUIRE(p->states_dim == 0 || b->states_buf != nullptr, "states_buf is required when states_dim > 0");
  HA_REQUIRE_ALIGNED16(b->dof_state);
  HA_REQUIRE_ALIGNED16(b->obs_buf);
  HA_REQUIRE_ALIGNED16(b->actions);
  if (b->dof_force) HA_REQUIRE_ALIGNED16(b->dof_force);
  if (b->force_sensors) HA_REQUIRE_ALIGNED16(b->force_sensors);
  if (b->states_buf) HA_REQUIRE_ALIGNED16(b->states_buf);
  if (p->root_regular)
    HA_REQUIRE(p->actors_per_env > 0 && p->object_slot >= 0 && p->object_slot < p->actors_per_env && p->goal_slot >= 0 &&
                   p->goal_slot < p->actors_per_env, "root_regular needs 0 <= object_slot, goal_slot < actors_per_env");
  return HA_OK;
}

int launch_hand_post(const HaHandParams* p, const HaHandBuffers* b, int mode, cudaStream_t st) {
  static ha::PerDeviceFlag flag;
  bool& configured = flag.here();
  const size_t smem = ha::hand_post_smem_bytes(p->n_dofs, p->n_actions, p->is_shadow ? p->n_tips : 0);
  if (!configured) {
    const size_t smem_max = ha::hand_post_smem_bytes(HA_HAND_MAX_DOFS, HA_HAND_MAX_DOFS, HA_HAND_MAX_TIPS);
    int rc = ha::check_cuda(cudaFuncSetAttribute(ha::hand_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_max),
                            "cudaFuncSetAttribute(hand_post_kernel)");
    if (rc) return rc;
    configured = true;
  }
  const int grid = (p->n_envs + ha::kHandTile - 1) / ha::kHandTile;
  return ha::check_cuda(ha::launch(ha::hand_post_kernel, grid, ha::kHandThreads, smem, st, *p, *b, mode), "hand_post_kernel launch");
}

}  // namespace

extern "C" {

size_t ha_sizeof_hand_params(void) { return sizeof(HaHandParams); }
size_t ha_sizeof_hand_buffers(void) { return sizeof(HaHandBuffers); }

int ha_hand_obs_dim(int is_shadow, int obs_type) {
  // shadow_hand.py:111-116, allegro_hand.py:108-112
  static const int shadow[4] = {42, 77, 157, 211};
  static const int allegro[4] = {-1, 50, 72, 88};
  if (obs_type < 0 || obs_type > 3) return -1;
  return is_shadow ? shadow[obs_type] : allegro[obs_type];
}

void ha_hand_fill_derived(HaHandParams* p) {
  if (!p) return;
  for (int d = 0; d < HA_HAND_MAX_DOFS; ++d) {
    const float range = p->dof_upper[d] - p->dof_lower[d];
    p->dof_range_rcp[d] = range != 0.0f ? (float)(1.0 / (double)range) : 0.0f;
  }
}

// SURVEY.md 8d counting rule: every needed input element once, every output element once, int64 surface buffers at 8 B
size_t ha_hand_post_bytes_per_env_step(const HaHandParams* p) {
  if (!p) return 0;
  const size_t D = (size_t)p->n_dofs, A = (size_t)p->n_actions, T = p->is_shadow ? (size_t)p->n_tips : 0;
  const bool full = p->obs_type == HA_HAND_OBS_FULL_STATE || p->states_dim > 0;
  const bool dof = p->obs_type != HA_HAND_OBS_OPENAI || p->states_dim > 0;
  size_t rd = 52 + 28 + 4 * A + 28;                    // object root row, goal pose, actions, reward state (3 int64 + successes)
  if (dof) rd += 8 * D;
  if (full) rd += 4 * D + 24 * T;
  rd += 52 * T;
  const size_t wr = 4 * ((size_t)p->obs_dim + (size_t)p->states_dim) + 4 + 24 + 4;       // rows, reward, 3 int64, successes
  return rd + wr;
}

size_t ha_hand_pre_bytes_per_env_step(const HaHandParams* p) {
  if (!p) return 0;
  const size_t A = (size_t)p->n_actions;
  return 5 * 4 * A + 16 + (p->force_enabled ? 28 : 0);  // actions r, prev r, cur w, prev w, actions copy w, 2 int64 masks, force row r+w + prob
}

size_t ha_hand_bytes_per_env_step(const HaHandParams* p) { return ha_hand_post_bytes_per_env_step(p) + ha_hand_pre_bytes_per_env_step(p); }

int ha_hand_pre_step(const HaHandParams* p, const HaHandBuffers* b, const float* actions, const float* ext_rand, ha_stream_t stream) {
  int rc = validate_hand(p, b);
  if (rc) return rc;
  HA_REQUIRE(actions != nullptr, "actions is null");
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  const int grid = (p->n_envs + ha::kEnvsPerCta - 1) / ha::kEnvsPerCta;
  if (grid < ha::kHandNPart) { rc = ha_hand_stats_finalize(p, b, stream); if (rc) return rc; }      // too few CTAs to fold a statistic each
  return ha::check_cuda(ha::launch(ha::hand_pre_kernel, grid, ha::kThreads, 0, (cudaStream_t)stream, *p, *b, actions, ext_rand,
                                   (int)ha::kHandPreFull), "hand_pre_kernel launch");
}

int ha_hand_reset_idx(const HaHandParams* p, const HaHandBuffers* b, const float* ext_rand, ha_stream_t stream) {
  int rc = validate_hand(p, b);
  if (rc) return rc;
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  const int grid = (p->n_envs + ha::kEnvsPerCta - 1) / ha::kEnvsPerCta;
  if (grid < ha::kHandNPart) { rc = ha_hand_stats_finalize(p, b, stream); if (rc) return rc; }
  rc = ha::check_cuda(ha::launch(ha::hand_pre_kernel, grid, ha::kThreads, 0, (cudaStream_t)stream, *p, *b, (const float*)nullptr, ext_rand,
                                 (int)ha::kHandPreResetOnly), "hand_pre_kernel(reset_only) launch");
  if (rc) return rc;
  if (p->use_fast_rng) ha::hand_rng_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(b->dynamic);
  return ha::check_cuda(cudaGetLastError(), "hand_rng_tick_kernel launch");
}

int ha_hand_post_step(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream) {
  int rc = validate_hand(p, b);
  if (rc) return rc;
  return launch_hand_post(p, b, ha::kHandPostFull, (cudaStream_t)stream);
}

int ha_hand_stats_finalize(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream) {
  int rc = validate_hand(p, b);
  if (rc) return rc;
  if ((p->n_envs + ha::kHandTile - 1) / ha::kHandTile > HA_HAND_STAT_SLOTS) return HA_OK;      // finished inside the post launch
  return ha::check_cuda(ha::launch(ha::hand_stats_finalize_kernel, (int)ha::kHandNPart, 128, 0, (cudaStream_t)stream, *b, (int)p->n_envs,
                                   p->av_factor), "hand_stats_finalize_kernel launch");
}

int ha_hand_observe(const HaHandParams* p, const HaHandBuffers* b, ha_stream_t stream) {
  int rc = validate_hand(p, b);
  if (rc) return rc;
  return launch_hand_post(p, b, ha::kHandPostObserveOnly, (cudaStream_t)stream);
}

}  // extern "C"
