// K_pre for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1355-1424 (pre_physics_step), including the sparse
// reset branch: reset_target_pose (:1191-1196) -> _reset_target
// (allegro_kuka_reorientation.py:104-128), reset_idx (:1246-1353),
// reset_object_pose (:1198-1224), set_actor_root_state_tensor_indexed
// (:1229-1244, here: compacted int32 actor rows for the simulator), the
// action -> DOF-target path (:1377-1397) and the random object forces
// (:1401-1414).  No nonzero()/unique()/len(): masks are consumed in place.
//
// Mapping: CTA = tile of 32 envs, 128 threads.
//   all threads : (env, dof) elementwise pass over the contiguous [32, 23]
//                 tiles of actions / prev_targets / cur_targets / dof_state;
//   warp 0      : per-env scalar reset work (goal pose, object pose, episode
//                 state, actor index lists);
//   warp 1      : random force on the object body row;
//   warps 2,3   : zero the 12 per-episode accumulators of reset envs.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"

namespace ha {

enum { kPreFull = 0, kPreResetOnly = 1 };

struct KukaRand {
  const float* ext;   // [n_envs, stride] or nullptr
  int stride;
  Rng rng;
  __device__ __forceinline__ float u(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.uniform(env, slot);
  }
  __device__ __forceinline__ float nrm(int env, int slot, int j) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot + j) : rng.normal(env, j);
  }
};

// random_force_prob = exp((log lo - log hi) * u + log hi)   (:1265-1269)
__device__ __forceinline__ float force_prob_from(const HaKukaParams& p, float u) {
  return expf(p.force_log_c0 * u + p.force_log_c1);
}

__global__ void __launch_bounds__(kThreads, 8)
kuka_pre_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b,
                const float* __restrict__ actions, const float* __restrict__ ext_rand, const int mode) {
  constexpr int D = 23;
  __shared__ unsigned char s_reset[kEnvsPerCta];
  __shared__ int s_last;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int env0 = blockIdx.x * kEnvsPerCta;
  const int n = p.n_envs;
  const int nvalid = min(kEnvsPerCta, n - env0);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;

  KukaRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_KUKA_RND_STRIDE(D);
  R.rng.seed = b.dynamic->rng_seed;
  R.rng.step = b.dynamic->rng_step;

  // every warp reads the tile's masks (coalesced int64)
  const bool rst = valid && (b.reset_buf[env] != 0);
  const bool reset_only = (mode == kPreResetOnly);   // VecTask.reset_idx / reset_done outside of a step
  const bool grst = valid && !reset_only && (b.reset_goal_buf[env] != 0);
  if (warp == 0) s_reset[lane] = rst ? 1 : 0;
  __syncthreads();

  // ---------------- all threads: DOF targets, elementwise -----------------
  {
    const size_t base = (size_t)env0 * D;
    const int total = nvalid * D;
    constexpr int kIters = (kEnvsPerCta * D + kThreads - 1) / kThreads;  // 6
    float a[kIters], prev[kIters];
#pragma unroll
    for (int i = 0; i < kIters; ++i) {
      const int f = tid + i * kThreads;
      if (f < total) { a[i] = reset_only ? 0.f : __ldg(actions + base + f); prev[i] = b.prev_targets[base + f]; }
    }
#pragma unroll
    for (int i = 0; i < kIters; ++i) {
      const int f = tid + i * kThreads;
      if (f >= total) continue;
      const int e = f / D, d = f - e * D;
      if (reset_only && !s_reset[e]) continue;
      const float lo = p.dof_lower[d], hi = p.dof_upper[d];
      float pv = prev[i];
      if (s_reset[e]) {
        // reset_idx DOF re-randomisation (:1271-1290)
        const int ge = env0 + e;
        const float dflt = p.dof_default[d];
        const float delta_max = hi - dflt, delta_min = lo - dflt;
        const float u_pos = R.u(ge, HA_KUKA_RND_DOF_POS + d);
        const float rand_delta = delta_min + (delta_max - delta_min) * u_pos;
        const float pos = dflt + p.dof_reset_noise[d] * rand_delta;
        const float u_vel = R.u(ge, HA_KUKA_RND_DOF_POS + D + d);
        const float vel = p.reset_dof_vel_noise * (2.0f * u_vel + -1.0f);
        *reinterpret_cast<float2*>(b.dof_state + 2 * (base + f)) = make_float2(pos, vel);
        pv = pos;   // prev_targets = cur_targets = allegro_pos (:1289-1290)
      }
      float cur;
      if (reset_only) {
        cur = pv;
      } else if (d >= 7) {   // absolute control of the hand DOFs (:1377-1390)
        cur = scale_to(a[i], lo, hi);
        cur = p.act_moving_average * cur + p.one_minus_act_moving_average * pv;
        cur = clamp_to(cur, lo, hi);
      } else {        // relative control of the arm DOFs (:1392-1395)
        cur = clamp_to(pv + p.arm_action_gain * a[i], lo, hi);
      }
      b.cur_targets[base + f] = cur;
      b.prev_targets[base + f] = cur;     // (:1397)
      if (!reset_only) b.actions[base + f] = a[i];   // self.actions = actions.clone() (:1357)
    }
  }

  // ---------------- warp 0: per-env scalar reset work ---------------------
  if (warp == 0) {
    const bool any_goal = rst || grst;
    if (any_goal) {
      // _reset_target; an env that is in both sets is drawn twice by the
      // reference and the reset_idx draw wins
      const int sp = rst ? HA_KUKA_RND_RESET_GOAL_POS : HA_KUKA_RND_GOAL_POS;
      const int sr = rst ? HA_KUKA_RND_RESET_GOAL_ROT : HA_KUKA_RND_GOAL_ROT;
      float g[7];
#pragma unroll
      for (int c = 0; c < 3; ++c) g[c] = p.target_min[c] + R.u(env, sp + c) * p.target_size[c];
      const f4 q = random_quat(R.u(env, sr + 0), R.u(env, sr + 1), R.u(env, sr + 2));
      g[3] = q.x; g[4] = q.y; g[5] = q.z; g[6] = q.w;
      float* gs = b.goal_states + (size_t)env * 13;
      const int grow = b.goal_actor[env];
      float* rr = b.root_state + (size_t)grow * 13;
#pragma unroll
      for (int c = 0; c < 7; ++c) { gs[c] = g[c]; rr[c] = g[c]; }
#pragma unroll
      for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
      b.reset_goal_buf[env] = 0;                       // reset_target_pose (:1194-1196)
      b.near_goal_steps[env] = 0;
      b.closest_keypoint_max_dist[env] = -1.0f;
    }
    if (rst) {
      // reset_object_pose (:1198-1224)
      const float* init = b.object_init_state + (size_t)env * 13;
      const int orow = b.object_actor[env];
      float* rr = b.root_state + (size_t)orow * 13;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        const float r = 2.0f * R.u(env, HA_KUKA_RND_OBJ_POS + c) + -1.0f;     // torch_rand_float(-1, 1)
        rr[c] = __ldg(init + c) + p.reset_pos_noise[c] * r;
      }
      const f4 q = random_quat(R.u(env, HA_KUKA_RND_OBJ_ROT + 0), R.u(env, HA_KUKA_RND_OBJ_ROT + 1), R.u(env, HA_KUKA_RND_OBJ_ROT + 2));
      rr[3] = q.x; rr[4] = q.y; rr[5] = q.z; rr[6] = q.w;
#pragma unroll
      for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
      *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(-1.f, -1.f, -1.f, -1.f);
      b.furthest_hand_dist[env] = -1.0f;
      b.random_force_prob[env] = force_prob_from(p, R.u(env, HA_KUKA_RND_FORCE_PROB));
      // episode bookkeeping (:1323-1347)
      b.progress_buf[env] = 0;
      b.reset_buf[env] = 0;
      b.prev_episode_successes[env] = b.successes[env];
      b.successes[env] = 0.0f;
      b.prev_episode_true_objective[env] = b.true_objective[env];
      b.true_objective[env] = 0.0f;
      b.lifted_object[env] = 0;
    }
    // actor rows for set_actor_root_state_tensor_indexed / set_dof_*_indexed:
    // warp-aggregated append, rows of one env stay adjacent and ascending
    const unsigned m_any = __ballot_sync(0xffffffffu, any_goal);
    const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
    const unsigned m_goal_only = m_any & ~m_rst;
    const int n_root = 2 * __popc(m_rst) + __popc(m_goal_only);
    const int n_hand = __popc(m_rst);
    int root_base = 0, hand_base = 0;
    if (lane == 0 && n_root) {
      root_base = atomicAdd(b.reset_counts + 0, n_root);
      if (n_hand) hand_base = atomicAdd(b.reset_counts + 1, n_hand);
      if (n_hand) atomicAdd(b.reset_counts + 2, n_hand);
      atomicAdd(b.reset_counts + 3, __popc(m_any));
    }
    root_base = __shfl_sync(0xffffffffu, root_base, 0);
    hand_base = __shfl_sync(0xffffffffu, hand_base, 0);
    const unsigned below = (1u << lane) - 1u;
    if (any_goal) {
      const int pos = root_base + 2 * __popc(m_rst & below) + __popc(m_goal_only & below);
      int k = pos;
      if (rst) b.root_reset_indices[k++] = b.object_actor[env];
      b.root_reset_indices[k] = b.goal_actor[env];
    }
    if (rst) b.hand_reset_indices[hand_base + __popc(m_rst & below)] = b.hand_actor[env];
  }

  // ---------------- warp 1: random forces on the object body ---------------
  if (warp == 1 && p.force_enabled && valid && !reset_only) {
    float* f = b.rb_forces + ((size_t)env * p.n_bodies + p.object_body) * 3;
    float fx = f[0], fy = f[1], fz = f[2];
    if (rst) { fx = 0.f; fy = 0.f; fz = 0.f; }          // rb_forces[env_ids] = 0 (:1255)
    fx *= p.force_decay_pow; fy *= p.force_decay_pow; fz *= p.force_decay_pow;   // (:1402)
    const float prob = rst ? force_prob_from(p, R.u(env, HA_KUKA_RND_FORCE_PROB)) : b.random_force_prob[env];
    if (R.u(env, HA_KUKA_RND_FORCE_TEST) < prob) {      // (:1405-1410)
      fx = R.nrm(env, HA_KUKA_RND_FORCE_NORMAL, 0) * p.force_gain[0] * p.force_gain[1];
      fy = R.nrm(env, HA_KUKA_RND_FORCE_NORMAL, 1) * p.force_gain[0] * p.force_gain[1];
      fz = R.nrm(env, HA_KUKA_RND_FORCE_NORMAL, 2) * p.force_gain[0] * p.force_gain[1];
    }
    f[0] = fx; f[1] = fy; f[2] = fz;
  }
  if (warp == 1 && (!p.force_enabled || reset_only) && rst) {
    float* f = b.rb_forces + ((size_t)env * p.n_bodies + p.object_body) * 3;
    f[0] = 0.f; f[1] = 0.f; f[2] = 0.f;
  }

  // ---------------- warps 2,3: zero per-episode accumulators ---------------
  if (warp >= 2 && rst) {
#pragma unroll
    for (int k = warp - 2; k < HA_KUKA_N_EPISODE_TERMS; k += 2) b.rewards_episode[(size_t)k * n + env] = 0.0f;   // (:1341-1342)
  }

  // ---------------- last CTA: advance the fast-mode RNG counter ------------
  if (p.use_fast_rng) {
    __threadfence();
    if (tid == 0) {
      const unsigned t = atomicAdd(b.tickets + 1, 1u);
      s_last = (t == gridDim.x - 1);
      if (s_last) { b.tickets[1] = 0; b.dynamic->rng_step = R.rng.step + 1; }
    }
  }
}

}  // namespace ha
