// K_pre for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1355-1424 (pre_physics_step), including the sparse
// reset branch: reset_target_pose (:1191-1196) -> _reset_target
// (allegro_kuka_reorientation.py:104-128), reset_idx (:1246-1353),
// reset_object_pose (:1198-1224), set_actor_root_state_tensor_indexed
// (:1229-1244, here: compacted int32 actor rows for the simulator), the
// action -> DOF-target path (:1377-1397) and the random object forces
// (:1401-1414).  No nonzero()/unique()/len(): masks are consumed in place.
//
// Mapping: see kuka_pre_kernel below.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"
#include "ha_kuka_post.cuh"   // kuka_stats_finalize_cta

namespace ha {

enum { kPreFull = 0, kPreResetOnly = 1 };

struct KukaRand {
  const float* ext;   // [n_envs, stride] or nullptr
  int stride;
  Rng rng;
  __device__ __noinline__ float u(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.uniform(env, slot);
  }
  // the four draws slot .. slot+3 (parity table: 4 loads; fast mode: ONE Philox block, id = slot)
  __device__ __noinline__ float4 u4(int env, int slot) const {
    if (ext) {
      const float* r = ext + (size_t)env * stride + slot;
      return make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
    }
    return rng.uniform4(env, 32 + slot);
  }
  __device__ __noinline__ float nrm(int env, int slot, int j) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot + j) : rng.normal(env, j);
  }
  // inlined single draw (the every-step Bernoulli test of the random forces)
  __device__ __forceinline__ float ui(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.uniform(env, slot);
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

// random_force_prob = exp((log lo - log hi) * u + log hi)   (:1265-1269)
__device__ __forceinline__ float force_prob_from(const HaKukaParams& p, float u) {
  return expf(p.force_log_c0 * u + p.force_log_c1);
}

struct KukaPreSmem {
  float2 lim[24];                  // (lower, upper) limit per DOF
  long long fold[kThreads / 32];   // statistics fold of the CTAs below kNPart (kuka_fold_*)
};

__device__ __forceinline__ float4 kuka_draw4(const KukaRand& R, int env, int slot) {   // inlined: independent Philox chains overlap
  if (R.ext) {
    const float* r = R.ext + (size_t)env * R.stride + slot;
    return make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
  }
  return R.rng.uniform4(env, 32 + slot);
}

// One actor's 13-float root state: a 52-byte row (Isaac Gym layout) or 13 plane elements n_envs apart (HA_STATE_SOA)
struct RootRow {
  float* base;
  size_t stride;
  __device__ __forceinline__ float& at(int c) const { return base[(size_t)c * stride]; }
};
__device__ __forceinline__ RootRow root_row(const HaKukaParams& p, const HaKukaBuffers& b, int actor_row, int env, int slot) {
  if (p.state_layout == HA_STATE_SOA) return {b.root_state + (size_t)slot * 13 * p.n_envs + env, (size_t)p.n_envs};
  return {b.root_state + (size_t)actor_row * 13, 1};
}
__device__ __forceinline__ int kuka_goal_row(const HaKukaParams& p, const HaKukaBuffers& b, int env) {
  return p.root_regular ? env * p.actors_per_env + p.goal_slot : b.goal_actor[env];
}
__device__ __forceinline__ int kuka_object_row(const HaKukaParams& p, const HaKukaBuffers& b, int env) {
  return p.root_regular ? env * p.actors_per_env + p.object_slot : b.object_actor[env];
}

// action -> DOF target of one element (:1377-1397): absolute control with a moving average for the hand DOFs, relative control
// for the 7 arm DOFs
__device__ __forceinline__ float kuka_dof_target(float av, float pv, int d, float2 lh, float ema, float one_m_ema, float gain) {
  float x;
  if (d >= 7) {
    x = scale_to(av, lh.x, lh.y);
    x = ema * x + one_m_ema * pv;
    x = clamp_to(x, lh.x, lh.y);
  } else {
    x = clamp_to(pv + gain * av, lh.x, lh.y);
  }
  return x;
}

// The whole re-randomisation of ONE env by ONE warp, one independent piece per lane: every piece starts with one Philox block (or
// its row of the parity table), so a single pass of the generator serves all of them and the dependent chain of the reset -- which
// bounds the launch, the tiles with a reset are its last ones to finish -- is one generator pass + half a random_quat + a few stores.
//   lanes 0..22  DOF d: reset_idx DOF re-randomisation (:1271-1290), prev_targets = cur_targets = pos (:1289-1290), then the
//                action pass on that position                                                          [full resets only]
//   lane 23      goal position, reset_target_pose (:1191-1196) with the subtask's _reset_target:
//                  reorientation  allegro_kuka_reorientation.py:104-128  position in the target volume (+ rotation: lanes 25, 26)
//                  regrasping     allegro_kuka_regrasping.py:76-100      position only (the goal ball keeps its rotation)
//                  throw          allegro_kuka_throw.py:90-117           bucket left / right of the table (4 draws), goal = bucket + 5 cm
//                an env that is in both sets is drawn twice by the reference and the reset_idx draw wins:
//                `full` selects the RESET_GOAL slots
//   lane 24      object position, reset_object_pose (:1198-1224) and its tail (:1221-1224) + lifted_object
//   lanes 25,26  goal rotation (reorientation): get_random_quat (:1178-1189) has two independent halves, (x, w) = sqrt(1 - u) *
//                (cos, sin)(2 pi v) and (y, z) = sqrt(u) * (sin, cos)(2 pi w) -- one lane each; goal row velocities = 0
//   lanes 27,28  object rotation, same split; object row velocities = 0
//                (regrasping / throw also put the object back on the table on a goal-only reset)
__device__ __forceinline__ void kuka_reset_env(const HaKukaParams& p, const HaKukaBuffers& b, const KukaRand R, const KukaPreSmem& s,
                                            const float* __restrict__ actions, const int env, const int lane, const bool full,
                                            const bool reset_only) {
  constexpr int D = 23;
  const bool reorient = p.subtask == HA_KUKA_REORIENTATION;
  const bool with_object = full || !reorient;
  // lane roles without branches: special lanes r = 0..5
  static_assert(HA_KUKA_RND_GOAL_POS == 0 && HA_KUKA_RND_GOAL_ROT == 3 && HA_KUKA_RND_RESET_GOAL_POS == 6 && HA_KUKA_RND_RESET_GOAL_ROT == 9 &&
                HA_KUKA_RND_OBJ_POS == 12 && HA_KUKA_RND_OBJ_ROT == 15, "slot arithmetic below");
  const int r = lane - D;
  const bool rot_lane = r >= 2 && r < 6, goal_lane = r == 0 || r == 2 || r == 3;
  const bool active = lane < D ? full : ((unsigned)r < 6u && (r == 0 || (goal_lane ? reorient : with_object)));
  const int slot = goal_lane ? (full ? 6 : 0) + (rot_lane ? 3 : 0) : (rot_lane ? 15 : 12);
  if (!active) return;
  // the two loads of the chain go first (object_init_state is cold: a DRAM round trip)
  float act_v = 0.f, init0 = 0.f, init1 = 0.f, init2 = 0.f;
  if (lane < D) {
    if (!reset_only) act_v = __ldg(actions + (size_t)env * D + lane);
  } else if (r == 1) {
    const float* init = b.object_init_state + (size_t)env * 13;
    init0 = __ldg(init + 0); init1 = __ldg(init + 1); init2 = __ldg(init + 2);
  }
  float4 u;
  if (R.ext) {
    const float* t = R.ext + (size_t)env * R.stride;
    if (lane < D) u = make_float4(__ldg(t + HA_KUKA_RND_DOF_POS + lane), __ldg(t + HA_KUKA_RND_DOF_POS + D + lane), 0.f, 0.f);
    else u = make_float4(__ldg(t + slot), __ldg(t + slot + 1), __ldg(t + slot + 2), __ldg(t + slot + 3));
  } else {
    u = R.rng.uniform4(env, lane < D ? 64 + lane : 32 + slot);   // DOF d: block 64 + d, (.x, .y) = (pos, vel) draws
  }
  HA_TRACE_POINT_IF(0, 4, threadIdx.x == 23);            // draws in hand
  if (lane < D) {
    const int d = lane;
    const size_t idx = (size_t)env * D + d;
    // (default position and noise coefficient: read here, not staged with the limits -- only this path needs them, and every
    // parameter line a kernel touches in its prologue is a constant-cache miss in EVERY tile)
    const float lo = s.lim[d].x, hi = s.lim[d].y, dflt = p.dof_default[d], noise = p.dof_reset_noise[d];
    const float delta_max = hi - dflt, delta_min = lo - dflt;
    const float rand_delta = delta_min + (delta_max - delta_min) * u.x;
    const float pos = dflt + noise * rand_delta;
    const float vel = p.reset_dof_vel_noise * (2.0f * u.y + -1.0f);
    *reinterpret_cast<float2*>(b.dof_state + 2 * idx) = make_float2(pos, vel);
    const float x = reset_only ? pos : kuka_dof_target(act_v, pos, d, s.lim[d], p.act_moving_average,
                                                         p.one_minus_act_moving_average, p.arm_action_gain);
    b.cur_targets[idx] = x;
    b.prev_targets[idx] = x;
    HA_TRACE_POINT_IF(0, 3, threadIdx.x == 0);           // DOF rows stored
    return;
  }
  const RootRow rr = goal_lane ? root_row(p, b, kuka_goal_row(p, b, env), env, p.goal_slot)
                               : root_row(p, b, kuka_object_row(p, b, env), env, p.object_slot);
  float* gs = b.goal_states + (size_t)env * 13;
  if (rot_lane) {
    // one half of get_random_quat (the arithmetic of random_quat, ha_device.cuh)
    const float two_pi = 6.2831854820251465f;  // float32(2*pi)
    const bool second = r & 1;                 // lanes 25 / 27: (x, w) from (u, v); lanes 26 / 28: (z, y) from (u, w)
    const float sq = second ? sqrtf(u.x) : sqrtf(1.0f - u.x);
    float sn, cs;
    sincosf(two_pi * (second ? u.z : u.y), &sn, &cs);
    const float q_cos = sq * cs, q_sin = sq * sn;          // (x, w) or (z, y)
    const int i_cos = second ? 5 : 3, i_sin = second ? 4 : 6;
    rr.at(i_cos) = q_cos; rr.at(i_sin) = q_sin;
    const int v0 = second ? 10 : 7;
    rr.at(v0) = 0.0f; rr.at(v0 + 1) = 0.0f; rr.at(v0 + 2) = 0.0f;
    if (goal_lane) { gs[i_cos] = q_cos; gs[i_sin] = q_sin; }
    HA_TRACE_POINT_IF(0, 11, threadIdx.x == 25);         // rotations stored
  } else if (r == 0) {
    float x, y, z, gz;
    if (p.subtask == HA_KUKA_THROW) {
      // torch_rand_float(lo, hi) = (hi - lo) * u + lo
      const float left_right = 2.0f * u.x + -1.0f;
      x = left_right > 0.0f ? 0.5f : -0.5f;
      const float sgn = left_right > 0.0f ? 1.0f : (left_right < 0.0f ? -1.0f : 0.0f);
      x = x + sgn * (0.4f * u.y + 0.0f);
      y = 1.7f * u.z + -1.0f;
      z = 1.0f * u.w + 0.0f;
      gz = z + 0.05f;
    } else {
      x = p.target_min[0] + u.x * p.target_size[0];
      y = p.target_min[1] + u.y * p.target_size[1];
      z = p.target_min[2] + u.z * p.target_size[2];
      gz = z;
    }
    rr.at(0) = x; rr.at(1) = y; rr.at(2) = z;
    gs[0] = x; gs[1] = y; gs[2] = gz;
    b.reset_goal_buf[env] = 0;
    b.near_goal_steps[env] = 0;
    b.closest_keypoint_max_dist[env] = -1.0f;
  } else {   // r == 1: object position; torch_rand_float(-1, 1) = 2 u - 1
    HA_TRACE_POINT_IF(0, 12, threadIdx.x == 24 && init0 != 123456.f);   // object_init_state arrived
    rr.at(0) = init0 + p.reset_pos_noise[0] * (2.0f * u.x + -1.0f);
    rr.at(1) = init1 + p.reset_pos_noise[1] * (2.0f * u.y + -1.0f);
    rr.at(2) = init2 + p.reset_pos_noise[2] * (2.0f * u.z + -1.0f);
    *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(-1.f, -1.f, -1.f, -1.f);
    b.furthest_hand_dist[env] = -1.0f;
    b.lifted_object[env] = 0;
  }
}

// CTA = tile of 32 envs, 4 warps.
//   element pass (all threads): the contiguous [32, 23] tiles of actions / prev_targets / cur_targets as 128-bit vectors, loads
//   issued before the masks are known, no branch on the masks in the tiles without a reset (9 of 10 at the benchmark's reset rate).
//   warp 1: random forces on the object body (every step), random_force_prob of reset envs.
//   Tiles with a reset (uniform branch): the re-randomisation runs FIRST, one warp per env with one independent piece per lane
//   (kuka_reset_env: inlined, so that the kernel parameters stay constant-bank operands -- behind a call they become generic
//   loads from the parameter space, a dependent chain of L2 round trips that used to dominate these tiles), then the element pass,
//   which leaves the rows of the full resets alone; what stays per tile with lane = env: warp 2 half of the accumulators, warp 3
//   episode bookkeeping + the other half + the actor-row lists (returning atomic issued right after the ballots, consumed last).
//   One CTA barrier, right after the ballots: it publishes the per-DOF tables and marks the point from which the masks may be
//   cleared (every warp holds its ballots).
__global__ void __launch_bounds__(kThreads, 8)
kuka_pre_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b,
                const float* __restrict__ actions, const float* __restrict__ ext_rand, const int mode) {
  constexpr int D = 23;
  constexpr int kVecPerTile = kEnvsPerCta * D / 4;     // 184 float4 per full tile
  constexpr int kIters = (kVecPerTile + kThreads - 1) / kThreads;   // 2
  __shared__ KukaPreSmem s;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int env0 = blockIdx.x * kEnvsPerCta;
  const int n = p.n_envs;
  const int nvalid = min(kEnvsPerCta, n - env0);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;
  const bool reset_only = (mode == kPreResetOnly);   // VecTask.reset_idx / reset_done outside of a step

  HA_TRACE_BEGIN(0);
  pdl_launch_dependents();
  pdl_wait();
  // ---- CTAs 0 .. kNPart-1 also fold one statistic of the previous post step each (loads in flight with the tile's) ----
  const bool fold = blockIdx.x < kNPart;                // uniform over the CTA
  KukaStatFold sf;
  if (fold) kuka_fold_load(b, blockIdx.x, tid, sf);
  // ---- issue the tile's vector loads first: they do not depend on the masks ----
  const size_t base = (size_t)env0 * D;
  const int total = nvalid * D, nvec = total >> 2;     // tile base is 16-byte aligned (env0 % 32 == 0)
  float4 a[kIters], prev[kIters];
#pragma unroll
  for (int i = 0; i < kIters; ++i) {
    const int q = tid + i * kThreads;
    a[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    prev[i] = a[i];
    if (q < nvec && !reset_only) {
      a[i] = __ldg(reinterpret_cast<const float4*>(actions + base) + q);
      prev[i] = *(reinterpret_cast<const float4*>(b.prev_targets + base) + q);
    }
  }
  // every warp reads the tile's masks (coalesced int64) -> identical ballots in all warps
  const long long raw_rst = valid ? b.reset_buf[env] : 0;
  const long long raw_goal = (valid && !reset_only) ? b.reset_goal_buf[env] : 0;
  // RNG state: device-resident (graph replays see the ticks)
  const unsigned long long rng_seed = b.dynamic->rng_seed, rng_step = b.dynamic->rng_step;
  // force inputs of warp 1 (issued early as well)
  const bool do_force = (warp == 1) && p.force_enabled && !reset_only;
  float f_prob = 0.f, f_v[3] = {0.f, 0.f, 0.f};
  const size_t frow = ((size_t)env0 * p.n_bodies + p.object_body) * 3;
  const int fstride = p.n_bodies * 3;
  // HA_STATE_SOA: rb_forces is [n_bodies, 3, n_envs] component planes like the state tensors -- lane = env, three coalesced rows
  const bool force_planes = p.state_layout == HA_STATE_SOA;
  float* const fplane = b.rb_forces + (size_t)p.object_body * 3 * n + env;
  if (do_force) {
    if (valid) f_prob = b.random_force_prob[env];
    if (force_planes) {
#pragma unroll
      for (int c = 0; c < 3; ++c)
        if (valid) f_v[c] = fplane[(size_t)c * n];
    } else {
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int idx = lane + 32 * i, e = idx / 3, c = idx - 3 * e;
        if (e < nvalid) f_v[i] = b.rb_forces[frow + (size_t)e * fstride + c];
      }
    }
  }
  if (tid < D) {
    s.lim[tid] = make_float2(p.dof_lower[tid], p.dof_upper[tid]);
  }
  HA_TRACE_POINT(0, 1);                                 // loads issued
  const bool rst = raw_rst != 0, grst = raw_goal != 0;
  const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
  const unsigned m_goal = __ballot_sync(0xffffffffu, grst);
  const unsigned m_any = m_rst | m_goal;               // uniform over the CTA
  HA_TRACE_POINT(0, 2);                                 // masks in
  // warp 3 of a tile with a reset: the list-slot atomic and the bookkeeping loads go in flight now and are consumed at the very end
  const unsigned m_goal_only = m_any & ~m_rst;
  // regrasping / throw: a new goal also puts the object back on the table (_reset_target of those subtasks), so a goal-only reset
  // touches the object row, the fingertip distances and lifted_object too and contributes two actor rows instead of one
  const bool goal_moves_object = p.subtask != HA_KUKA_REORIENTATION;
  const int rows_goal_only = goal_moves_object ? 2 : 1;
  unsigned long long list_old = 0;
  if (warp == 3 && m_any) {
    // actor rows for set_actor_root_state_tensor_indexed / set_dof_*_indexed: warp-aggregated append, rows of one env stay adjacent.
    // The four counters share one L2 line: one returning 64-bit add for the two list bases and one fire-and-forget add for the
    // two plain counts (reset_counts is 8-byte aligned; < 2^31 each)
    if (lane == 0) {
      const int n_root = 2 * __popc(m_rst) + rows_goal_only * __popc(m_goal_only), n_hand = __popc(m_rst);
      unsigned long long* rc = reinterpret_cast<unsigned long long*>(b.reset_counts);
      list_old = atomicAdd(rc + 0, (unsigned long long)(unsigned)n_root | ((unsigned long long)(unsigned)n_hand << 32));
      atomicAdd(rc + 1, (unsigned long long)(unsigned)n_hand | ((unsigned long long)(unsigned)__popc(m_any) << 32));
    }
  }
  if (fold) kuka_fold_warp(sf, blockIdx.x, tid, s.fold);
  __syncthreads();       // tables and fold partials visible; every warp holds its ballots: the masks may be cleared from here on
  HA_TRACE_POINT(0, 5);                                 // barrier
  if (fold && tid == kThreads - 1) kuka_fold_finish(b, blockIdx.x, n, s.fold);
  KukaRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_KUKA_RND_STRIDE(D);
  R.rng.seed = rng_seed;
  R.rng.step = rng_step;

  // ---------------- all threads: DOF targets, 4 consecutive elements per float4 ----------------
  const float ema = p.act_moving_average, one_m_ema = p.one_minus_act_moving_average, gain = p.arm_action_gain;
  const bool copy_actions = (actions != b.actions);
  // `skip`: envs of the tile whose rows kuka_reset_env writes (full reset); 0 in all but a few tiles -> plain 128-bit stores
  auto element_pass = [&](const unsigned skip) {
#pragma unroll
    for (int i = 0; i < kIters; ++i) {
      const int q = tid + i * kThreads;
      if (q >= nvec) continue;
      const int f0 = 4 * q;
      const int e0 = f0 / D;
      const int d0 = f0 - e0 * D;
      int d = d0;
      float4 o;
      o.x = kuka_dof_target(a[i].x, prev[i].x, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      o.y = kuka_dof_target(a[i].y, prev[i].y, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      o.z = kuka_dof_target(a[i].z, prev[i].z, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      o.w = kuka_dof_target(a[i].w, prev[i].w, d, s.lim[d], ema, one_m_ema, gain);
      if (skip == 0u) {
        *(reinterpret_cast<float4*>(b.cur_targets + base) + q) = o;
        *(reinterpret_cast<float4*>(b.prev_targets + base) + q) = o;       // (:1397)
      } else {
        const float ov[4] = {o.x, o.y, o.z, o.w};
        int e = e0, dd = d0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (!((skip >> e) & 1u)) {
            b.cur_targets[base + f0 + j] = ov[j];
            b.prev_targets[base + f0 + j] = ov[j];
          }
          if (++dd == D) { dd = 0; ++e; }
        }
      }
      if (copy_actions) *(reinterpret_cast<float4*>(b.actions + base) + q) = a[i];   // self.actions = actions.clone() (:1357)
    }
    // ragged last tile: up to 3 trailing elements
    const int f = (nvec << 2) + tid;
    if (f < total) {
      const int e = f / D, d = f - e * D;
      const float av = __ldg(actions + base + f);
      if (!((skip >> e) & 1u)) {
        const float x = kuka_dof_target(av, b.prev_targets[base + f], d, s.lim[d], ema, one_m_ema, gain);
        b.cur_targets[base + f] = x;
        b.prev_targets[base + f] = x;
      }
      if (copy_actions) b.actions[base + f] = av;
    }
  };
  // ---------------- warp 1: random forces on the object body, random_force_prob of reset envs.  FIRST: its inputs arrive with the
  // masks, its generator pass then runs while the tile's vectors are still in flight ----------------
  if (warp == 1) {
    float prob = f_prob;
    if (rst) {                                           // (:1265-1269)
      prob = force_prob_from(p, kuka_draw4(R, env, HA_KUKA_RND_FORCE_PROB).x);
      b.random_force_prob[env] = prob;
    }
    if (do_force) {
      // Bernoulli test per env (lane = env), then the force row is decayed / kicked (:1401-1414)
      bool hit = false;
      if (valid) hit = R.ui(env, HA_KUKA_RND_FORCE_TEST) < prob;
      const unsigned m_hit = __ballot_sync(0xffffffffu, hit);
      const float decay = p.force_decay_pow, g0 = p.force_gain[0], g1 = p.force_gain[1];
      // the three normals of a kicked env: its row of the parity table, or ONE counter block in fast mode (lane = env)
      float nrm[3] = {0.f, 0.f, 0.f};
      if (hit) R.nrm3(env, HA_KUKA_RND_FORCE_NORMAL, nrm);
      if (force_planes) {
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          if (valid) {
            float v = rst ? 0.f : f_v[c];                          // rb_forces[env_ids] = 0 (:1255)
            v *= decay;                                            // (:1402)
            if (hit) v = nrm[c] * g0 * g1;                         // (:1405-1410)
            fplane[(size_t)c * n] = v;
          }
        }
      } else {
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int idx = lane + 32 * i, e = idx / 3, c = idx - 3 * e;
          float kick = 0.f;
          if (m_hit != 0u) {                                       // (uniform) component c of env e lives in lane e
            const float k0 = __shfl_sync(0xffffffffu, nrm[0], e & 31), k1 = __shfl_sync(0xffffffffu, nrm[1], e & 31), k2 = __shfl_sync(0xffffffffu, nrm[2], e & 31);
            kick = c == 0 ? k0 : (c == 1 ? k1 : k2);
          }
          if (e < nvalid) {
            float v = ((m_rst >> e) & 1u) ? 0.f : f_v[i];          // rb_forces[env_ids] = 0 (:1255)
            v *= decay;                                            // (:1402)
            if ((m_hit >> e) & 1u) v = kick * g0 * g1;             // (:1405-1410)
            b.rb_forces[frow + (size_t)e * fstride + c] = v;
          }
        }
      }
    } else if (rst) {
      if (force_planes) {
        fplane[0] = 0.f; fplane[(size_t)n] = 0.f; fplane[(size_t)2 * n] = 0.f;
      } else {
        float* f = b.rb_forces + frow + (size_t)lane * fstride;
        f[0] = 0.f; f[1] = 0.f; f[2] = 0.f;
      }
    }
  }

  // The same pass by the warps that carry no reset env (a tile with one, two or three of them: envs of rank 0, 1, 2 go to warps 0, 2,
  // 3, so warps {1, 2, 3}, {1, 3} or {1} are free): `n_free` warps x 32 lanes stride the 184 vectors of the tile, which come back
  // from L1 / L2 under the new thread mapping.  The launch ends with its slowest tile, and that is one with several resets.
  auto element_pass_free = [&](const unsigned skip, const int free_idx, const int n_free) {
    const int t = free_idx * 32 + lane, step = 32 * n_free;
    for (int q = t; q < nvec; q += step) {
      const float4 av = __ldg(reinterpret_cast<const float4*>(actions + base) + q);
      const float4 pv = *(reinterpret_cast<const float4*>(b.prev_targets + base) + q);
      const int f0 = 4 * q;
      const int e0 = f0 / D;
      const int d0 = f0 - e0 * D;
      int d = d0;
      float ov[4];
      ov[0] = kuka_dof_target(av.x, pv.x, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      ov[1] = kuka_dof_target(av.y, pv.y, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      ov[2] = kuka_dof_target(av.z, pv.z, d, s.lim[d], ema, one_m_ema, gain); d = (d + 1 == D) ? 0 : d + 1;
      ov[3] = kuka_dof_target(av.w, pv.w, d, s.lim[d], ema, one_m_ema, gain);
      int e = e0, dd = d0;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (!((skip >> e) & 1u)) {
          b.cur_targets[base + f0 + j] = ov[j];
          b.prev_targets[base + f0 + j] = ov[j];
        }
        if (++dd == D) { dd = 0; ++e; }
      }
      if (copy_actions) *(reinterpret_cast<float4*>(b.actions + base) + q) = av;
    }
    const int f = (nvec << 2) + t;                       // ragged last tile: up to 3 trailing elements
    if (f < total) {
      const int e = f / D, d = f - e * D;
      const float av = __ldg(actions + base + f);
      if (!((skip >> e) & 1u)) {
        const float x = kuka_dof_target(av, b.prev_targets[base + f], d, s.lim[d], ema, one_m_ema, gain);
        b.cur_targets[base + f] = x;
        b.prev_targets[base + f] = x;
      }
      if (copy_actions) b.actions[base + f] = av;
    }
  };
  if (m_any == 0u) {
    if (!reset_only) element_pass(0u);
    HA_TRACE_POINT(0, 6);                               // element pass stored
  } else {
    // tiles with a reset (uniform over the CTA): the re-randomisation FIRST -- it is the longest dependent chain of the launch --
    // one warp per env (kuka_reset_env); then the element pass, which leaves the rows of the full resets alone, so the two need no
    // ordering.
    float succ = 0.f, tobj = 0.f;
    int hand_row = 0;
    if (warp == 3 && rst) {                              // bookkeeping inputs: in flight under the re-randomisation
      hand_row = b.hand_actor[env];
      succ = b.successes[env];
      tobj = b.true_objective[env];
    }
    unsigned m = m_any;
    int rank = 0;
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1;
      // envs of rank 0, 1, 2, 3 -> warps 0, 2, 3, 1 (warp 1 also carries the forces, warp 3 the lists)
      if (((0x1320 >> ((rank++ & 3) * 4)) & 3) == warp) kuka_reset_env(p, b, R, s, actions, env0 + e, lane, (m_rst >> e) & 1u, reset_only);
    }
    HA_TRACE_POINT(0, 10);                              // warp 0: its envs are done
    if (!reset_only) {
      const int n_env = __popc(m_any);
      if (n_env > 3) element_pass(m_rst);                              // every warp carries an env: every warp keeps its share
      else {
        // busy warps: 0 | 0, 2 | 0, 2, 3  ->  free warps: 1, 2, 3 | 1, 3 | 1
        const int free_idx = n_env == 1 ? warp - 1 : (n_env == 2 ? (warp == 1 ? 0 : (warp == 3 ? 1 : -1)) : (warp == 1 ? 0 : -1));
        if (free_idx >= 0) element_pass_free(m_rst, free_idx, 4 - n_env);
      }
    }
    HA_TRACE_POINT(0, 6);                               // element pass stored
    // what stays per tile (lane = env): accumulators, episode bookkeeping, actor-row lists
    if (warp == 2) {
      if (rst) {
#pragma unroll
        for (int k = 0; k < HA_KUKA_N_EPISODE_TERMS; k += 2) b.rewards_episode[(size_t)k * n + env] = 0.0f;   // (:1341-1342)
      }
    } else if (warp == 3) {
      if (rst) {
        // episode bookkeeping (:1323-1347)
        b.progress_buf[env] = 0;
        b.reset_buf[env] = 0;
        b.prev_episode_successes[env] = succ;
        b.successes[env] = 0.0f;
        b.prev_episode_true_objective[env] = tobj;
        b.true_objective[env] = 0.0f;
#pragma unroll
        for (int k = 1; k < HA_KUKA_N_EPISODE_TERMS; k += 2) b.rewards_episode[(size_t)k * n + env] = 0.0f;
      }
      HA_TRACE_POINT_IF(0, 7, tid == 96);
      // list slots: the atomic issued before the barrier is consumed here, as the last thing this warp does
      const unsigned long long old = __shfl_sync(0xffffffffu, list_old, 0);
      const int root_base = (int)(unsigned)(old & 0xffffffffULL), hand_base = (int)(unsigned)(old >> 32);
      if (rst || grst) {
        const unsigned below = (1u << lane) - 1u;
        int k = root_base + 2 * __popc(m_rst & below) + rows_goal_only * __popc(m_goal_only & below);
        if (rst || goal_moves_object) b.root_reset_indices[k++] = kuka_object_row(p, b, env);
        if (rst) b.hand_reset_indices[hand_base + __popc(m_rst & below)] = hand_row;
        b.root_reset_indices[k] = kuka_goal_row(p, b, env);
      }
      HA_TRACE_POINT_IF(0, 8, tid == 96);
    }
  }

  HA_TRACE_POINT(0, 9);
#ifdef HA_TRACE
  if (tid == 0 && blockIdx.x < kTraceMaxCtas) g_ha_trace[0][blockIdx.x].c[13] = (m_any ? (m_rst ? 2 : 1) : 0) + 16 * __popc(m_any);   // kind + 16 x number of envs with a reset
#endif
  HA_TRACE_END(0);
}

// advances the fast-mode RNG counter after an out-of-step reset (inside a step ha_kuka_post_step does it)
__global__ void kuka_rng_tick_kernel(HaKukaDynamic* dyn) { dyn->rng_step = dyn->rng_step + 1; }

}  // namespace ha
