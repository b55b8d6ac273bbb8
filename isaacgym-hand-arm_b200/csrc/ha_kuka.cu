// C-ABI entry points of the AllegroKuka family (see include/handarm_b200.h).
#include <cuda_runtime.h>
#include <stdlib.h>

#include "../../include/handarm_b200.h"
#include "ha_host.h"
#include "ha_kuka_post.cuh"
#include "ha_kuka_pre.cuh"
#include "ha_kuka2_post.cuh"
#include "ha_kuka2_pre.cuh"

#ifdef HA_TRACE
// diagnostic builds only (not declared in the header): copies the records of the last pre / post launch to the host
extern "C" int ha_debug_trace_read(void* dst, size_t bytes) {
  if (bytes > sizeof(ha::HaTraceRec) * 2 * ha::kTraceMaxCtas) bytes = sizeof(ha::HaTraceRec) * 2 * ha::kTraceMaxCtas;
  return ha::check_cuda(cudaMemcpyFromSymbol(dst, ha::g_ha_trace, bytes), "cudaMemcpyFromSymbol(g_ha_trace)");
}
#endif

namespace {

int validate_kuka(const HaKukaParams* p, const HaKukaBuffers* b) {
  HA_REQUIRE(p != nullptr && b != nullptr, "null params / buffers");
  HA_REQUIRE(p->abi_version == HA_ABI_VERSION, "HaKukaParams.abi_version %d != %d", p->abi_version, HA_ABI_VERSION);
  HA_REQUIRE(p->n_envs > 0, "n_envs must be positive (got %d)", p->n_envs);
  HA_REQUIRE(p->n_arms == 1 || p->n_arms == 2, "n_arms must be 1 or 2 (got %d)", p->n_arms);
  const int A = p->n_arms;
  HA_REQUIRE(p->n_dofs == 23 * A && p->dofs_per_env == 23 * A, "AllegroKuka with %d arm(s) needs n_dofs == dofs_per_env == %d (got %d/%d)",
             A, 23 * A, p->n_dofs, p->dofs_per_env);
  HA_REQUIRE(p->n_tips == 4 * A, "n_tips must be %d (got %d)", 4 * A, p->n_tips);
  if (A == 2) {
    HA_REQUIRE(p->state_layout == HA_STATE_ISAACGYM, "the two-arm kernels read Isaac Gym's AoS state tensors (HA_STATE_ISAACGYM)");
    HA_REQUIRE(p->palm_body[1] >= 0 && p->palm_body[1] < p->n_bodies, "palm_body[1] out of range");
  }
  HA_REQUIRE(p->n_keypoints == 1 || p->n_keypoints == 4, "n_keypoints must be 1 or 4 (got %d)", p->n_keypoints);
  HA_REQUIRE(p->obs_dim == ha_kuka_obs_dim(p->n_arms, p->n_keypoints), "obs_dim %d does not match the task layout %d", p->obs_dim,
             ha_kuka_obs_dim(p->n_arms, p->n_keypoints));
  HA_REQUIRE(p->n_bodies > 0 && p->object_body >= 0 && p->object_body < p->n_bodies, "object_body out of range");
  HA_REQUIRE(p->palm_body[0] >= 0 && p->palm_body[0] < p->n_bodies, "palm_body out of range");
  for (int t = 0; t < p->n_tips; ++t)
    HA_REQUIRE(p->tip_body[t] >= 0 && p->tip_body[t] < p->n_bodies, "tip_body[%d] out of range", t);
  HA_REQUIRE(p->success_steps >= 1, "success_steps must be >= 1");
  HA_REQUIRE(p->post_grid >= 0, "post_grid must be >= 0 (got %d)", p->post_grid);
  HA_REQUIRE(p->subtask == HA_KUKA_REORIENTATION || p->subtask == HA_KUKA_REGRASPING || p->subtask == HA_KUKA_THROW, "unknown subtask %d",
             p->subtask);
  HA_REQUIRE(b->xchg == nullptr || (b->xchg_reduced != nullptr && p->n_arms == 1),
             "HaKukaBuffers.xchg needs xchg_reduced and the single-arm task (the in-kernel exchange lives in its post kernel)");
  const void* need[] = {b->root_state, b->rigid_body_state, b->dof_state, b->object_actor, b->goal_actor, b->hand_actor,
                        b->object_init_state, b->object_keypoint_offsets, b->object_scales, b->goal_states,
                        b->obs_buf, b->rew_buf, b->reset_buf, b->reset_goal_buf, b->progress_buf, b->timeout_buf, b->successes,
                        b->prev_episode_successes, b->true_objective, b->prev_episode_true_objective, b->near_goal_steps,
                        b->lifted_object, b->closest_keypoint_max_dist, b->closest_fingertip_dist,
                        b->rewards_episode, b->episode_cumulative, b->prev_targets, b->cur_targets, b->actions, b->rb_forces,
                        b->random_force_prob, b->root_reset_indices, b->hand_reset_indices, b->reset_counts, b->stats, b->stat_accum,
                        b->tickets, b->dynamic};
  for (size_t i = 0; i < sizeof(need) / sizeof(need[0]); ++i) HA_REQUIRE(need[i] != nullptr, "HaKukaBuffers field #%zu is null", i);
  HA_REQUIRE(A == 2 || (b->finger_rew_coeffs != nullptr && b->furthest_hand_dist != nullptr),
             "finger_rew_coeffs / furthest_hand_dist are required by the single-arm task");
  HA_REQUIRE_ALIGNED16(b->dof_state);
  HA_REQUIRE(((uintptr_t)b->reset_counts & 7) == 0, "reset_counts must be 8-byte aligned");
  HA_REQUIRE_ALIGNED16(b->obs_buf);
  HA_REQUIRE_ALIGNED16(b->closest_fingertip_dist);
  if (b->finger_rew_coeffs) HA_REQUIRE_ALIGNED16(b->finger_rew_coeffs);
  HA_REQUIRE_ALIGNED16(b->object_keypoint_offsets);
  HA_REQUIRE_ALIGNED16(b->object_scales);
  if (A == 1) HA_REQUIRE_ALIGNED16(b->goal_states);
  HA_REQUIRE_ALIGNED16(b->prev_targets);
  HA_REQUIRE_ALIGNED16(b->cur_targets);
  HA_REQUIRE_ALIGNED16(b->actions);
  HA_REQUIRE(p->state_layout == HA_STATE_ISAACGYM || p->state_layout == HA_STATE_SOA, "unknown state_layout %d", p->state_layout);
  HA_REQUIRE(p->state_layout != HA_STATE_SOA || p->root_regular, "HA_STATE_SOA needs the regular actor numbering (root_regular)");
  if (p->state_layout == HA_STATE_SOA) {
    HA_REQUIRE_ALIGNED16(b->root_state);
    HA_REQUIRE_ALIGNED16(b->rigid_body_state);
  }
  if (p->root_regular)
    HA_REQUIRE(p->actors_per_env > 0 && p->object_slot >= 0 && p->object_slot < p->actors_per_env && p->goal_slot < p->actors_per_env,
               "root_regular needs 0 <= object_slot, goal_slot < actors_per_env");
  return HA_OK;
}

// CTAs of the persistent post grid: HaKukaParams.post_grid (test / tuning hook, per call) wins over the environment
// variable HA_POST_GRID (A/B runs of a whole process), which wins over one full wave of resident CTAs.
int post_grid_limit(const HaKukaParams* p, int blocks_per_sm) {
  static int env_grid = -1;
  if (env_grid < 0) {
    const char* e = getenv("HA_POST_GRID");
    env_grid = e ? atoi(e) : 0;
    if (env_grid < 0) env_grid = 0;
  }
  if (p->post_grid > 0) return p->post_grid;
  return env_grid > 0 ? env_grid : ha::device_sm_count() * blocks_per_sm;
}

#ifndef HA_POST_NS
#define HA_POST_NS 2
#endif
constexpr int kPostNS = HA_POST_NS;   // cp.async stages per CTA (1: seven single-tile CTAs per SM; 2: four double-buffered CTAs)

template <int K, bool SOA>
int launch_post(const HaKukaParams* p, const HaKukaBuffers* b, int mode, cudaStream_t st) {
  using S = ha::KukaPostSmem<K, kPostNS>;
  static int per_device[ha::kMaxDevices] = {};
  int& blocks_per_sm = per_device[ha::current_device()];
  if (blocks_per_sm == 0) {
    int rc = ha::check_cuda(cudaFuncSetAttribute(ha::kuka_post_kernel<K, SOA, kPostNS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)),
                            "cudaFuncSetAttribute(kuka_post_kernel)");
    if (rc) return rc;
    int nb = 0;
    rc = ha::check_cuda(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, ha::kuka_post_kernel<K, SOA, kPostNS>, ha::kPostThreads, sizeof(S)),
                        "cudaOccupancyMaxActiveBlocksPerMultiprocessor(kuka_post_kernel)");
    if (rc) return rc;
    blocks_per_sm = nb > 0 ? nb : 1;
  }
  // persistent CTAs: one wave of resident CTAs walks the 32-env tiles
  const int n_tiles = (p->n_envs + ha::kTile - 1) / ha::kTile;
  int limit = post_grid_limit(p, blocks_per_sm);
  if (limit > HA_KUKA_STAT_SLOTS) limit = HA_KUKA_STAT_SLOTS;        // one statistics slot per CTA
  const int grid = n_tiles < limit ? n_tiles : limit;
  return ha::check_cuda(ha::launch(ha::kuka_post_kernel<K, SOA, kPostNS>, grid, ha::kPostThreads, sizeof(S), st, *p, *b, mode),
                        "kuka_post_kernel launch");
}

template <int K>
int launch_post2(const HaKukaParams* p, const HaKukaBuffers* b, int mode, cudaStream_t st) {
  using S = ha::Kuka2PostSmem<K>;
  static ha::PerDeviceFlag flag;
  bool& configured = flag.here();
  if (!configured) {
    int rc = ha::check_cuda(cudaFuncSetAttribute(ha::kuka2_post_kernel<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(S)),
                            "cudaFuncSetAttribute(kuka2_post_kernel)");
    if (rc) return rc;
    configured = true;
  }
  const int grid = (p->n_envs + ha::kTile - 1) / ha::kTile;
  return ha::check_cuda(ha::launch(ha::kuka2_post_kernel<K>, grid, ha::kPostThreads, sizeof(S), st, *p, *b, mode), "kuka2_post_kernel launch");
}

}  // namespace

extern "C" {

int ha_kuka_obs_dim(int n_arms, int n_keypoints) {
  if (n_arms == 1) return ha::kuka_layout(23, 4, n_keypoints).dim;
  // allegro_kuka_two_arms.py:176-211: no closest_fingertip_dist slots, keypoints relative to both palms
  if (n_arms == 2) return 46 + 46 + 6 + 20 + 10 + 24 + 6 * n_keypoints + 3 * n_keypoints + 3 + 1 + 1 + 1 + 1 + 1;
  return -1;
}

size_t ha_kuka_post_bytes_per_env_step(const HaKukaParams* p) {
  if (!p) return 0;
  const size_t D = (size_t)p->n_dofs, T = (size_t)p->n_tips, K = (size_t)p->n_keypoints, A = (size_t)p->n_arms;
  // reads: dof pos+vel, object root row, goal pose, palm rows, fingertip pos+rot, per-env constants, state, accumulators
  size_t rd = 8 * D + 52 + 28 + 52 * A + 28 * T + (12 * K + 12 + 4 + (A == 1 ? 4 * T : 0));
  rd += 4 + 4 * T + (A == 1 ? 4 : 0) + 1 + 4 + 4 + 8 + 8 + 4;
  const size_t n_ep = (A == 1) ? HA_KUKA_N_EPISODE_TERMS : 8, n_step = (A == 1) ? HA_KUKA_N_STEP_TERMS : 5;
  rd += 4 * n_ep;
  // writes: obs row, reward, 3 int64 flags, per-env state, accumulators, per-step terms
  size_t wr = 4 * (size_t)p->obs_dim + 4 + 24 + (4 + 4 * T + (A == 1 ? 4 : 0) + 1 + 4 + 4 + 4) + 4 * n_ep + 4 * n_step;
  return rd + wr;
}

size_t ha_kuka_pre_bytes_per_env_step(const HaKukaParams* p) {
  if (!p) return 0;
  const size_t D = (size_t)p->n_dofs;
  // actions r, prev_targets r, cur_targets w, prev_targets w, actions copy w, 2 int64 masks r, force row r+w, force prob r
  return 5 * 4 * D + 16 + 24 + 4;
}

size_t ha_kuka_bytes_per_env_step(const HaKukaParams* p) {
  return ha_kuka_post_bytes_per_env_step(p) + ha_kuka_pre_bytes_per_env_step(p);
}

void ha_kuka_fill_derived(HaKukaParams* p) {
  if (!p) return;
  for (int d = 0; d < HA_MAX_DOFS; ++d) {
    const float range = p->dof_upper[d] - p->dof_lower[d];          // fp32 difference, as the kernel forms it
    p->dof_range_rcp[d] = range != 0.0f ? (float)(1.0 / (double)range) : 0.0f;
  }
}

int64_t ha_kuka_stat_accum_init(int i) { return ha::stat_is_max(i) ? (int64_t)ha::kStatMaxInit : 0; }

int ha_kuka_pre_step(const HaKukaParams* p, const HaKukaBuffers* b, const float* actions, const float* ext_rand, ha_stream_t stream) {
  int rc = validate_kuka(p, b);
  if (rc) return rc;
  HA_REQUIRE(actions != nullptr, "actions is null");
  HA_REQUIRE(ha::aligned16(actions), "actions must be 16-byte aligned");
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  const int grid = (p->n_envs + ha::kEnvsPerCta - 1) / ha::kEnvsPerCta;
  if (grid < ha::kNPart) { rc = ha_kuka_stats_finalize(p, b, stream); if (rc) return rc; }      // too few CTAs to fold a statistic each
  if (p->n_arms == 2) {
    HA_REQUIRE(p->subtask == HA_KUKA_REGRASPING || p->subtask == HA_KUKA_REORIENTATION, "two arms: regrasping or reorientation (subtask %d)", p->subtask);
    return ha::check_cuda(ha::launch(ha::kuka2_pre_kernel, grid, ha::kThreads, 0, (cudaStream_t)stream, *p, *b, actions, ext_rand,
                                     (int)ha::kPreFull), "kuka2_pre_kernel launch");
  }
  return ha::check_cuda(ha::launch(ha::kuka_pre_kernel, grid, ha::kThreads, 0, (cudaStream_t)stream, *p, *b, actions, ext_rand, (int)ha::kPreFull),
                        "kuka_pre_kernel launch");
}

int ha_kuka_reset_idx(const HaKukaParams* p, const HaKukaBuffers* b, const float* ext_rand, ha_stream_t stream) {
  int rc = validate_kuka(p, b);
  if (rc) return rc;
  HA_REQUIRE(p->use_fast_rng || ext_rand != nullptr, "ext_rand is required unless use_fast_rng is set");
  const int grid = (p->n_envs + ha::kEnvsPerCta - 1) / ha::kEnvsPerCta;
  if (grid < ha::kNPart) { rc = ha_kuka_stats_finalize(p, b, stream); if (rc) return rc; }      // too few CTAs to fold a statistic each
  if (p->n_arms == 2) {
    HA_REQUIRE(p->subtask == HA_KUKA_REGRASPING || p->subtask == HA_KUKA_REORIENTATION, "two arms: regrasping or reorientation (subtask %d)", p->subtask);
    ha::kuka2_pre_kernel<<<grid, ha::kThreads, 0, (cudaStream_t)stream>>>(*p, *b, nullptr, ext_rand, ha::kPreResetOnly);
  } else {
    ha::kuka_pre_kernel<<<grid, ha::kThreads, 0, (cudaStream_t)stream>>>(*p, *b, nullptr, ext_rand, ha::kPreResetOnly);
  }
  if (p->use_fast_rng) ha::kuka_rng_tick_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(b->dynamic);
  return ha::check_cuda(cudaGetLastError(), "kuka_pre_kernel(reset_only) launch");
}

static int post_dispatch(const HaKukaParams* p, const HaKukaBuffers* b, int mode, ha_stream_t stream) {
  int rc = validate_kuka(p, b);
  if (rc) return rc;
  if (p->n_arms == 2) return p->n_keypoints == 4 ? launch_post2<4>(p, b, mode, (cudaStream_t)stream) : launch_post2<1>(p, b, mode, (cudaStream_t)stream);
  const bool soa = p->state_layout == HA_STATE_SOA;
  if (p->n_keypoints == 4) return soa ? launch_post<4, true>(p, b, mode, (cudaStream_t)stream) : launch_post<4, false>(p, b, mode, (cudaStream_t)stream);
  return soa ? launch_post<1, true>(p, b, mode, (cudaStream_t)stream) : launch_post<1, false>(p, b, mode, (cudaStream_t)stream);
}

int ha_kuka_post_step(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream) {
  return post_dispatch(p, b, ha::kPostFull, stream);
}

int ha_kuka_observe(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream) {
  return post_dispatch(p, b, ha::kPostObserveOnly, stream);
}

int ha_kuka_stats_finalize(const HaKukaParams* p, const HaKukaBuffers* b, ha_stream_t stream) {
  int rc = validate_kuka(p, b);
  if (rc) return rc;
  if (p->n_arms == 2 && (p->n_envs + ha::kTile - 1) / ha::kTile > HA_KUKA_STAT_SLOTS) return HA_OK;      // finalised inside the post launch
  return ha::check_cuda(ha::launch(ha::kuka_stats_finalize_kernel, (int)ha::kNPart, ha::kThreads, 0, (cudaStream_t)stream, *b, p->n_envs), "kuka_stats_finalize_kernel launch");
}

}  // extern "C"
