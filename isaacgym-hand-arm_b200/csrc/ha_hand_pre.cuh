// K_pre for the in-hand reorientation family (ShadowHand, AllegroHand): one launch replaces pre_physics_step
// (shadow_hand.py:670-708 / allegro_hand.py:586-625) including the sparse reset branch: reset_target_pose
// (:586-602 / :506-522), reset_idx (:604-668 / :524-584), randomize_rotation (:803-806), the compacted actor rows
// for set_actor_root_state_tensor_indexed / set_dof_*_indexed, the action -> DOF-target path through
// actuated_dof_indices and the random object forces.  Masks are consumed in place (no nonzero / unique / len).
//
// Mapping: CTA = 32-env tile, 128 threads.  All threads: the [32, n_actions] action tile, flat (coalesced), each
// element scattered to its actuated DOF column; warp 0: per-env scalar reset work (lane = env); reset envs get
// their DOFs re-randomised by one warp per env (lane = DOF) through shared memory; warp 1: random forces.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"
#include "ha_hand_post.cuh"      // hand_fold_*: the statistics fold of the previous post step

namespace ha {

enum { kHandPreFull = 0, kHandPreResetOnly = 1 };

struct HandRand {
  const float* ext;   // [n_envs, stride] or nullptr (fast mode)
  int stride;
  Rng rng;
  __device__ __forceinline__ float u(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.uniform(env, slot);
  }
  __device__ __forceinline__ float nrm(int env, int slot, int j) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot + j) : rng.normal(env, j);
  }
  // the three normals slot .. slot+2 of one env: its row of the parity table, or ONE counter block in fast mode
  __device__ __forceinline__ void nrm3(int env, int slot, float (&o)[3]) const {
    if (ext) {
      const float* r = ext + (size_t)env * stride + slot;
      o[0] = __ldg(r); o[1] = __ldg(r + 1); o[2] = __ldg(r + 2);
    } else {
      rng.normal3(env, (uint32_t)slot, o);
    }
  }
};

// quat_from_angle_axis(angle, unit axis e_k) (torch_jit_utils.py:118-123): normalize([axis * sin(a/2), cos(a/2)])
__device__ __forceinline__ f4 quat_axis_angle(float angle, int axis) {
  const float th = angle / 2.0f;
  const float s = sinf(th), c = cosf(th);
  // normalize(axis) = axis / clamp(|axis|, 1e-9) = axis for a unit axis; quat_unit divides by the 4-norm
  const float nrm = fmaxf(sqrtf(s * s + c * c), 1e-9f);
  f4 q = {0.f, 0.f, 0.f, c / nrm};
  const float v = s / nrm;
  if (axis == 0) q.x = v; else if (axis == 1) q.y = v; else q.z = v;
  return q;
}
// randomize_rotation (shadow_hand.py:803-806): quat_mul(aa(r0 * pi, x), aa(r1 * pi, y))
__device__ __forceinline__ f4 hand_random_rotation(float r0, float r1) {
  const float pi = 3.1415927410125732f;   // float32(np.pi)
  return quat_mul(quat_axis_angle(r0 * pi, 0), quat_axis_angle(r1 * pi, 1));
}

// randomize_rotation_pen (shadow_hand.py:809-814, max_angle = 0.3): quat_mul(aa(pi / 2 + r0 * 0.3, x), aa(r0 * pi, z))
__device__ __forceinline__ f4 hand_random_rotation_pen(float r0) {
  const float pi = 3.1415927410125732f, half_pi = 1.5707963705062866f;   // float32(np.pi), float32(0.5 * np.pi)
  return quat_mul(quat_axis_angle(half_pi + r0 * 0.3f, 0), quat_axis_angle(r0 * pi, 2));
}

struct HandPreSmem {
  float pos[kEnvsPerCta][HA_HAND_MAX_DOFS];   // re-randomised DOF positions of the tile's reset envs
  long long fold[8];                          // statistics fold of the CTAs below kHandNPart (hand_fold_*)
};

__global__ void __launch_bounds__(kThreads)
hand_pre_kernel(const __grid_constant__ HaHandParams p, const __grid_constant__ HaHandBuffers b, const float* __restrict__ actions,
                const float* __restrict__ ext_rand, const int mode) {
  __shared__ HandPreSmem s;
  const int D = p.n_dofs, A = p.n_actions, Dp = p.dofs_per_env;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.n_envs;
  const int env0 = blockIdx.x * kEnvsPerCta;
  const int nvalid = min(kEnvsPerCta, n - env0);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;
  const bool reset_only = (mode == kHandPreResetOnly);

  // CTAs 0 .. kHandNPart-1 also fold one statistic of the previous post step each (loads in flight with the tile's)
  const bool fold = blockIdx.x < kHandNPart && (n + kHandTile - 1) / kHandTile <= HA_HAND_STAT_SLOTS;      // uniform over the CTA
  HandStatFold sf;
  if (fold) hand_fold_load(reinterpret_cast<const long long*>(b.stat_accum), blockIdx.x, tid, sf);

  HandRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_HAND_RND_STRIDE(D);
  R.rng.seed = b.dynamic->rng_seed;
  R.rng.step = b.dynamic->rng_step;

  // every warp reads the tile's masks (coalesced int64) -> identical ballots in all warps
  const bool rst = valid && b.reset_buf[env] != 0;
  const bool grst = valid && !reset_only && b.reset_goal_buf[env] != 0;
  const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
  const unsigned m_goal = __ballot_sync(0xffffffffu, grst);
  const unsigned m_any = m_rst | m_goal;

  // ---------------- DOF re-randomisation of reset envs: one warp per env, lane = DOF (:632-645 / :560-573) ----------------
  if (m_rst != 0u) {
    unsigned m = m_rst;
    int rank = 0;
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1;
      if ((rank++ & (kThreads / 32 - 1)) != warp || lane >= D) continue;
      const int d = lane, re = env0 + e;
      const float r_pos = 2.0f * R.u(re, HA_HAND_RND_RESET + 5 + d) + -1.0f;
      const float r_vel = 2.0f * R.u(re, HA_HAND_RND_RESET + 5 + D + d) + -1.0f;
      const float lo = p.dof_lower[d], hi = p.dof_upper[d], dflt = p.dof_default[d];
      const float delta_max = hi - dflt, delta_min = lo - dflt;
      const float rand_delta = delta_min + (delta_max - delta_min) * 0.5f * (r_pos + 1.0f);
      const float pos = dflt + p.reset_dof_pos_noise * rand_delta;
      const float vel = 0.0f + p.reset_dof_vel_noise * r_vel;
      *reinterpret_cast<float2*>(b.dof_state + 2 * ((size_t)re * Dp + d)) = make_float2(pos, vel);
      s.pos[e][d] = pos;
      b.prev_targets[(size_t)re * Dp + d] = pos;
      b.cur_targets[(size_t)re * Dp + d] = pos;
    }
  }
  if (fold) hand_fold_warp(sf, tid, s.fold);
  __syncthreads();
  if (fold && tid == kThreads - 1) hand_fold_finish(b.stats, b.consecutive_successes, b.tickets, blockIdx.x, n, p.av_factor, s.fold);

  // ---------------- all threads: actions -> DOF targets of the actuated DOFs ----------------
  if (!reset_only) {
    const bool copy_actions = (actions != b.actions);
    const int total = nvalid * A;
    const float* a_src = actions + (size_t)env0 * A;
    for (int idx = tid; idx < total; idx += kThreads) {
      const int e = idx / A, a = idx - e * A;
      const int d = p.actuated_dof[a];
      const float av = __ldg(a_src + idx);
      const size_t t = (size_t)(env0 + e) * Dp + d;
      const float prev = ((m_rst >> e) & 1u) ? s.pos[e][d] : b.prev_targets[t];
      const float lo = p.dof_lower[d], hi = p.dof_upper[d];
      float x;
      if (p.use_relative_control) {                  // (:685-688)
        x = clamp_to(prev + p.rel_action_gain * av, lo, hi);
      } else {                                       // (:689-695)
        x = scale_to(av, lo, hi);
        x = p.act_moving_average * x + p.one_minus_act_moving_average * prev;
        x = clamp_to(x, lo, hi);
      }
      b.cur_targets[t] = x;
      b.prev_targets[t] = x;                         // (:697)
      if (copy_actions) b.actions[(size_t)env0 * A + idx] = av;   // self.actions = actions.clone() (:684)
    }
  }

  // ---------------- warp 1: random forces on the object body (:699-708) ----------------
  if (warp == 1) {
    float* frow = b.rb_forces + ((size_t)env * p.n_bodies + p.object_body) * 3;
    if (p.force_enabled && !reset_only) {
      if (valid) {
        float prob = b.random_force_prob[env];
        if (rst) prob = expf(p.force_log_c0 * R.u(env, HA_HAND_RND_FORCE_PROB) + p.force_log_c1);   // drawn in reset_idx, used below
        float f[3] = {0.f, 0.f, 0.f};                // rb_forces[env_ids] = 0 (:614)
        if (!rst) { f[0] = frow[0]; f[1] = frow[1]; f[2] = frow[2]; }          // in flight under the generator pass
        const bool hit = R.u(env, HA_HAND_RND_FORCE_TEST) < prob;
        float kick[3] = {0.f, 0.f, 0.f};
        if (hit) R.nrm3(env, HA_HAND_RND_FORCE_NORMAL, kick);                  // one counter block in fast mode
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          float v = f[c];
          v *= p.force_decay_pow;
          if (hit) v = kick[c] * p.force_gain[0] * p.force_gain[1];
          frow[c] = v;
        }
      }
    } else if (rst) {
      frow[0] = 0.f; frow[1] = 0.f; frow[2] = 0.f;
    }
  }

  // ---------------- warp 0: per-env scalar reset work, lane = env ----------------
  if (warp != 0 || m_any == 0u) return;
  const unsigned m_goal_only = m_any & ~m_rst;
  const int n_root = 2 * __popc(m_rst) + __popc(m_goal_only), n_hand = __popc(m_rst);
  int root_base = 0, hand_base = 0;
  if (lane == 0) {
    root_base = atomicAdd(b.reset_counts + 0, n_root);
    if (n_hand) { hand_base = atomicAdd(b.reset_counts + 1, n_hand); atomicAdd(b.reset_counts + 2, n_hand); }
    atomicAdd(b.reset_counts + 3, __popc(m_any));
  }
  root_base = __shfl_sync(0xffffffffu, root_base, 0);
  hand_base = __shfl_sync(0xffffffffu, hand_base, 0);
  const bool any_goal = rst || grst;
  if (!any_goal) return;
  const int goal_row = p.root_regular ? env * p.actors_per_env + p.goal_slot : b.goal_actor[env];
  {
    // reset_target_pose (:586-602); an env in both sets is drawn twice by the reference and the reset_idx draw wins
    const int slot = rst ? HA_HAND_RND_RESET_GOAL : HA_HAND_RND_GOAL;
    const float r0 = 2.0f * R.u(env, slot + 0) + -1.0f, r1 = 2.0f * R.u(env, slot + 1) + -1.0f;
    const f4 q = hand_random_rotation(r0, r1);
    const float* gi = b.goal_init_state + (size_t)env * 13;
    float* gs = b.goal_states + (size_t)env * 13;
    float* rr = b.root_state + (size_t)goal_row * 13;
#pragma unroll
    for (int c = 0; c < 3; ++c) { const float g = __ldg(gi + c); gs[c] = g; rr[c] = g + p.goal_displacement[c]; }
    gs[3] = q.x; gs[4] = q.y; gs[5] = q.z; gs[6] = q.w;
    rr[3] = q.x; rr[4] = q.y; rr[5] = q.z; rr[6] = q.w;
#pragma unroll
    for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
    b.reset_goal_buf[env] = 0;
  }
  const unsigned below = (1u << lane) - 1u;
  int k = root_base + 2 * __popc(m_rst & below) + __popc(m_goal_only & below);
  if (rst) {
    // reset_idx: object pose (:616-631 / :533-548)
    const int obj_row = p.root_regular ? env * p.actors_per_env + p.object_slot : b.object_actor[env];
    const float* init = b.object_init_state + (size_t)env * 13;
    float* rr = b.root_state + (size_t)obj_row * 13;
    float r[5];
#pragma unroll
    for (int c = 0; c < 5; ++c) r[c] = 2.0f * R.u(env, HA_HAND_RND_RESET + c) + -1.0f;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
      const float v0 = __ldg(init + c);
      rr[c] = (c < 2 || c == p.up_axis_idx) ? v0 + p.reset_position_noise * r[c] : v0;
    }
    const f4 q = p.ignore_z_rot ? hand_random_rotation_pen(r[3]) : hand_random_rotation(r[3], r[4]);   // objectType "pen" (:626-629)
    rr[3] = q.x; rr[4] = q.y; rr[5] = q.z; rr[6] = q.w;
#pragma unroll
    for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
    b.random_force_prob[env] = expf(p.force_log_c0 * R.u(env, HA_HAND_RND_FORCE_PROB) + p.force_log_c1);   // (:641-643)
    b.progress_buf[env] = 0;
    b.reset_buf[env] = 0;
    b.successes[env] = 0.0f;
    b.root_reset_indices[k++] = obj_row;
    b.hand_reset_indices[hand_base + __popc(m_rst & below)] = b.hand_actor[env];
  }
  b.root_reset_indices[k] = goal_row;
}

__global__ void hand_rng_tick_kernel(HaHandDynamic* dyn) { dyn->rng_step = dyn->rng_step + 1; }

}  // namespace ha
