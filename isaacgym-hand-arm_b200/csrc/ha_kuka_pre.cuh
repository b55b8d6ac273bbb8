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
//   all threads : the contiguous [32, 23] tiles of actions / prev_targets /
//                 cur_targets as 128-bit vectors (one float4 = 4 consecutive
//                 (env, dof) elements); loads are issued before the masks are
//                 known, the masks arrive as two 32-bit ballots;
//   warp 0      : per-env scalar reset work, only for tiles that have a reset; its loads,
//                 random draws and list-slot atomics are issued as soon as the masks are
//                 known, so a tile with a reset pays ONE extra memory round trip;
//   warp 1      : random force on the object body rows, 3 coalesced requests;
//   warps 2,3   : zero the 12 per-episode accumulators of reset envs.
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
};

// random_force_prob = exp((log lo - log hi) * u + log hi)   (:1265-1269)
__device__ __forceinline__ float force_prob_from(const HaKukaParams& p, float u) {
  return expf(p.force_log_c0 * u + p.force_log_c1);
}

constexpr int kSpecDraws = 5;      // Philox blocks of a full reset drawn ahead of the masks (kuka_spec_slot)
struct KukaPreSmem {
  float lo[24], hi[24], dflt[24], noise[24];
  float pos[kEnvsPerCta][24];      // re-randomised DOF positions of the tile's reset envs (row = env in tile)
  long long fold[kThreads / 32];   // statistics fold of the CTAs below kNPart (kuka_fold_*)
  float4 spec[kSpecDraws][kEnvsPerCta];      // fast-RNG draws of the scalar reset, made ahead of the masks (kuka_spec_slot)
};

// reset_idx DOF re-randomisation of one env by one warp, lane = dof (:1271-1290).
// Writes dof_state; the new position goes to smem for the element pass (prev_targets = cur_targets = pos, :1289-1290),
// and straight to global memory when there is no element pass (reset outside of a step).
__device__ __noinline__ void kuka_reset_dofs_warp(const HaKukaBuffers& b, const KukaRand& R, KukaPreSmem& s, float vel_noise,
                                                  int env, int e_local, int lane, bool write_targets, const HaKukaParams& p) {
  constexpr int D = 23;
  if (lane >= D) return;
  const int d = lane;
  float u_pos, u_vel;
  if (R.ext) {
    u_pos = __ldg(R.ext + (size_t)env * R.stride + HA_KUKA_RND_DOF_POS + d);
    u_vel = __ldg(R.ext + (size_t)env * R.stride + HA_KUKA_RND_DOF_POS + D + d);
  } else {
    const float2 u = R.rng.uniform2(env, 64 + d);     // one Philox block per DOF: (.x, .y) = (pos, vel) draws
    u_pos = u.x; u_vel = u.y;
  }
  // limits straight from the kernel parameters: this runs ahead of the barrier that publishes the shared copies
  const float lo = p.dof_lower[d], hi = p.dof_upper[d], dflt = p.dof_default[d], noise = p.dof_reset_noise[d];
  const float delta_max = hi - dflt, delta_min = lo - dflt;
  const float rand_delta = delta_min + (delta_max - delta_min) * u_pos;
  const float pos = dflt + noise * rand_delta;
  const float vel = vel_noise * (2.0f * u_vel + -1.0f);
  const size_t idx = (size_t)env * D + d;
  *reinterpret_cast<float2*>(b.dof_state + 2 * idx) = make_float2(pos, vel);
  s.pos[e_local][d] = pos;
  if (write_targets) { b.prev_targets[idx] = pos; b.cur_targets[idx] = pos; }
}

// Everything the scalar reset of one env needs, fetched / drawn up front (lane = env).
struct KukaResetPrefetch {
  float4 u_gp, u_gr, u_op, u_or;   // goal position / rotation, object position / rotation draws
  float u_fp, succ, tobj, init0, init1, init2;
  int goal_row, obj_row, hand_row, root_pos, hand_pos, root_base, hand_base;
};

__device__ __forceinline__ float4 kuka_draw4(const KukaRand& R, int env, int slot) {   // inlined: independent Philox chains overlap
  if (R.ext) {
    const float* r = R.ext + (size_t)env * R.stride + slot;
    return make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
  }
  return R.rng.uniform4(env, 32 + slot);
}

// Fast-RNG mode: the Philox draws of the scalar reset do not depend on the masks (only the choice between the two goal slots
// does), so they are computed while the mask loads are still in flight (~1 us of DRAM latency in which the warps would
// otherwise sit at the ballot) instead of on the critical path of the tiles that turn out to have a reset: warp w draws blocks
// w and w + 4 of the five of a full reset for its lane's env into shared memory, warp 0 picks them up after the CTA barrier.
// (a goal-only reset draws its two blocks once the masks are known)
__device__ __forceinline__ int kuka_spec_slot(int k) {
  switch (k) {
    case 0: return HA_KUKA_RND_RESET_GOAL_POS;
    case 1: return HA_KUKA_RND_RESET_GOAL_ROT;
    case 2: return HA_KUKA_RND_OBJ_POS;
    case 3: return HA_KUKA_RND_OBJ_ROT;
    default: return HA_KUKA_RND_FORCE_PROB;
  }
}

// Issue phase: loads, random draws and the list-slot atomics of a tile that has a goal / full reset.
__device__ __forceinline__ void kuka_reset_prefetch(const HaKukaParams& p, const HaKukaBuffers& b, const KukaRand& R, int env, int lane,
                                                bool rst, bool grst, KukaResetPrefetch& f) {
  const bool any_goal = rst || grst;
  // actor rows for set_actor_root_state_tensor_indexed / set_dof_*_indexed:
  // warp-aggregated append, rows of one env stay adjacent and ascending
  const unsigned m_any = __ballot_sync(0xffffffffu, any_goal);
  const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
  const unsigned m_goal_only = m_any & ~m_rst;
  const int n_root = 2 * __popc(m_rst) + __popc(m_goal_only);
  const int n_hand = __popc(m_rst);
  int root_base = 0, hand_base = 0;
  if (lane == 0 && n_root) {
    // a third of the CTAs arrive here at the same moment and the four counters share one L2 line: one returning 64-bit add
    // for the two list bases and one fire-and-forget add for the two plain counts (reset_counts is 8-byte aligned; < 2^31 each)
    unsigned long long* rc = reinterpret_cast<unsigned long long*>(b.reset_counts);
    const unsigned long long old = atomicAdd(rc + 0, (unsigned long long)(unsigned)n_root | ((unsigned long long)(unsigned)n_hand << 32));
    root_base = (int)(unsigned)(old & 0xffffffffULL);
    hand_base = (int)(unsigned)(old >> 32);
    atomicAdd(rc + 1, (unsigned long long)(unsigned)n_hand | ((unsigned long long)(unsigned)__popc(m_any) << 32));
  }
  // _reset_target; an env that is in both sets is drawn twice by the
  // reference and the reset_idx draw wins
  const int sp = rst ? HA_KUKA_RND_RESET_GOAL_POS : HA_KUKA_RND_GOAL_POS;
  const int sr = rst ? HA_KUKA_RND_RESET_GOAL_ROT : HA_KUKA_RND_GOAL_ROT;
  f.succ = 0.f; f.tobj = 0.f; f.init0 = 0.f; f.init1 = 0.f; f.init2 = 0.f;
  f.goal_row = 0; f.obj_row = 0; f.hand_row = 0;
  if (any_goal) f.goal_row = p.root_regular ? env * p.actors_per_env + p.goal_slot : b.goal_actor[env];
  if (rst) {
    f.obj_row = p.root_regular ? env * p.actors_per_env + p.object_slot : b.object_actor[env];
    f.hand_row = b.hand_actor[env];
    f.succ = b.successes[env];
    f.tobj = b.true_objective[env];
    const float* init = b.object_init_state + (size_t)env * 13;
    f.init0 = __ldg(init + 0); f.init1 = __ldg(init + 1); f.init2 = __ldg(init + 2);
  }
  if (R.ext) {
    const int ec = any_goal ? env : 0;                // idle lanes read the row of env 0 (keeps the warp converged)
    f.u_gp = kuka_draw4(R, ec, sp);
    f.u_gr = kuka_draw4(R, ec, sr);
    f.u_op = kuka_draw4(R, ec, HA_KUKA_RND_OBJ_POS);
    f.u_or = kuka_draw4(R, ec, HA_KUKA_RND_OBJ_ROT);
    f.u_fp = kuka_draw4(R, ec, HA_KUKA_RND_FORCE_PROB).x;
  } else if (m_goal_only) {                           // fast mode: the full-reset blocks were drawn ahead of the masks (kuka_spec_slot)
    f.u_gp = kuka_draw4(R, env, HA_KUKA_RND_GOAL_POS);
    f.u_gr = kuka_draw4(R, env, HA_KUKA_RND_GOAL_ROT);
  }
  const unsigned below = (1u << lane) - 1u;
  f.root_pos = 2 * __popc(m_rst & below) + __popc(m_goal_only & below);   // offsets inside the warp's list segments;
  f.hand_pos = __popc(m_rst & below);                                     // the bases (lane 0's atomics) are added on use
  f.root_base = root_base;
  f.hand_base = hand_base;
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

// Apply phase: goal / object / episode-state reset of one env (lane = env) from the prefetched values.
__device__ __forceinline__ void kuka_reset_apply(const HaKukaParams& p, const HaKukaBuffers& b, const KukaResetPrefetch& f, int env,
                                                 bool rst, bool grst) {
  const bool any_goal = rst || grst;
  if (any_goal) {
    float g[7];
    g[0] = p.target_min[0] + f.u_gp.x * p.target_size[0];
    g[1] = p.target_min[1] + f.u_gp.y * p.target_size[1];
    g[2] = p.target_min[2] + f.u_gp.z * p.target_size[2];
    const f4 q = random_quat(f.u_gr.x, f.u_gr.y, f.u_gr.z);
    g[3] = q.x; g[4] = q.y; g[5] = q.z; g[6] = q.w;
    float* gs = b.goal_states + (size_t)env * 13;
    const RootRow rr = root_row(p, b, f.goal_row, env, p.goal_slot);
#pragma unroll
    for (int c = 0; c < 7; ++c) { gs[c] = g[c]; rr.at(c) = g[c]; }
#pragma unroll
    for (int c = 7; c < 13; ++c) rr.at(c) = 0.0f;
    b.reset_goal_buf[env] = 0;                       // reset_target_pose (:1194-1196)
    b.near_goal_steps[env] = 0;
    b.closest_keypoint_max_dist[env] = -1.0f;
  }
  if (rst) {
    // reset_object_pose (:1198-1224); torch_rand_float(-1, 1) = 2 u - 1
    const RootRow rr = root_row(p, b, f.obj_row, env, p.object_slot);
    rr.at(0) = f.init0 + p.reset_pos_noise[0] * (2.0f * f.u_op.x + -1.0f);
    rr.at(1) = f.init1 + p.reset_pos_noise[1] * (2.0f * f.u_op.y + -1.0f);
    rr.at(2) = f.init2 + p.reset_pos_noise[2] * (2.0f * f.u_op.z + -1.0f);
    const f4 q = random_quat(f.u_or.x, f.u_or.y, f.u_or.z);
    rr.at(3) = q.x; rr.at(4) = q.y; rr.at(5) = q.z; rr.at(6) = q.w;
#pragma unroll
    for (int c = 7; c < 13; ++c) rr.at(c) = 0.0f;
    *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(-1.f, -1.f, -1.f, -1.f);
    b.furthest_hand_dist[env] = -1.0f;
    b.random_force_prob[env] = force_prob_from(p, f.u_fp);
    // episode bookkeeping (:1323-1347)
    b.progress_buf[env] = 0;
    b.reset_buf[env] = 0;
    b.prev_episode_successes[env] = f.succ;
    b.successes[env] = 0.0f;
    b.prev_episode_true_objective[env] = f.tobj;
    b.true_objective[env] = 0.0f;
    b.lifted_object[env] = 0;
  }
  const int root_base = __shfl_sync(0xffffffffu, f.root_base, 0);
  const int hand_base = __shfl_sync(0xffffffffu, f.hand_base, 0);
  if (any_goal) {
    int k = root_base + f.root_pos;
    if (rst) b.root_reset_indices[k++] = f.obj_row;
    b.root_reset_indices[k] = f.goal_row;
  }
  if (rst) b.hand_reset_indices[hand_base + f.hand_pos] = f.hand_row;
}

__global__ void __launch_bounds__(kThreads, 7)
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
  // force inputs of warp 1 (issued early as well)
  const bool do_force = (warp == 1) && p.force_enabled && !reset_only;
  float f_prob = 0.f, f_v[3] = {0.f, 0.f, 0.f};
  const size_t frow = ((size_t)env0 * p.n_bodies + p.object_body) * 3;
  const int fstride = p.n_bodies * 3;
  if (do_force) {
    if (valid) f_prob = b.random_force_prob[env];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int idx = lane + 32 * i, e = idx / 3, c = idx - 3 * e;
      if (e < nvalid) f_v[i] = b.rb_forces[frow + (size_t)e * fstride + c];
    }
  }
  KukaRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_KUKA_RND_STRIDE(D);
  R.rng.seed = b.dynamic->rng_seed;
  R.rng.step = b.dynamic->rng_step;
  if (tid < D) {
    s.lo[tid] = p.dof_lower[tid]; s.hi[tid] = p.dof_upper[tid];
    s.dflt[tid] = p.dof_default[tid]; s.noise[tid] = p.dof_reset_noise[tid];
  }
  if (!R.ext) {
#pragma unroll
    for (int k = 0; k < 2; ++k)
      if (warp + 4 * k < kSpecDraws) s.spec[warp + 4 * k][lane] = R.rng.uniform4(env, 32 + kuka_spec_slot(warp + 4 * k));
  }
  const bool rst = raw_rst != 0, grst = raw_goal != 0;
  const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
  const unsigned m_goal = __ballot_sync(0xffffffffu, grst);
  const bool tile_resets = (m_rst | m_goal) != 0u;     // uniform over the CTA; about a third of the tiles per step
  // warp 0: everything the scalar reset needs goes in flight now and is consumed at the very end
  KukaResetPrefetch pf;
  if (warp == 0 && tile_resets) kuka_reset_prefetch(p, b, R, env, lane, rst, grst, pf);
  if (fold) kuka_fold_warp(sf, blockIdx.x, tid, s.fold);
  if (m_rst != 0u) {
    // DOF re-randomisation of the tile's reset envs, one warp per env (round robin), lane = dof; before the barrier, so that it
    // overlaps warp 0's atomics and the barrier below also publishes s.pos to the element pass
    unsigned m = m_rst;
    int rank = 0;
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1;
      if ((rank++ & (kThreads / 32 - 1)) == warp)
        kuka_reset_dofs_warp(b, R, s, p.reset_dof_vel_noise, env0 + e, e, lane, reset_only, p);
    }
  }
  __syncthreads();
  if (fold && tid == kThreads - 1) kuka_fold_finish(b, blockIdx.x, n, s.fold);
  if (warp == 0 && tile_resets && !R.ext) {
    if (rst) { pf.u_gp = s.spec[0][lane]; pf.u_gr = s.spec[1][lane]; }
    pf.u_op = s.spec[2][lane]; pf.u_or = s.spec[3][lane]; pf.u_fp = s.spec[4][lane].x;
  }

  // ---------------- all threads: DOF targets, 4 consecutive elements per float4 ----------------
  if (!reset_only) {
    const float ema = p.act_moving_average, one_m_ema = p.one_minus_act_moving_average, gain = p.arm_action_gain;
    const bool copy_actions = (actions != b.actions);
#pragma unroll
    for (int i = 0; i < kIters; ++i) {
      const int q = tid + i * kThreads;
      if (q >= nvec) continue;
      const int f0 = 4 * q;
      int e = f0 / D, d = f0 - e * D;
      const float av[4] = {a[i].x, a[i].y, a[i].z, a[i].w};
      const float pvv[4] = {prev[i].x, prev[i].y, prev[i].z, prev[i].w};
      float cur[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float lo = s.lo[d], hi = s.hi[d];
        float pv = pvv[c];
        if ((m_rst >> e) & 1u) pv = s.pos[e][d];       // re-randomised just above
        float x;
        if (d >= 7) {          // absolute control of the hand DOFs (:1377-1390)
          x = scale_to(av[c], lo, hi);
          x = ema * x + one_m_ema * pv;
          x = clamp_to(x, lo, hi);
        } else {               // relative control of the arm DOFs (:1392-1395)
          x = clamp_to(pv + gain * av[c], lo, hi);
        }
        cur[c] = x;
        if (++d == D) { d = 0; ++e; }
      }
      const float4 o = make_float4(cur[0], cur[1], cur[2], cur[3]);
      *(reinterpret_cast<float4*>(b.cur_targets + base) + q) = o;
      *(reinterpret_cast<float4*>(b.prev_targets + base) + q) = o;       // (:1397)
      if (copy_actions) *(reinterpret_cast<float4*>(b.actions + base) + q) = a[i];   // self.actions = actions.clone() (:1357)
    }
    // ragged last tile: up to 3 trailing elements
    const int f = (nvec << 2) + tid;
    if (f < total) {
      const int e = f / D, d = f - e * D;
      const float lo = s.lo[d], hi = s.hi[d];
      const float av = __ldg(actions + base + f);
      const float pv = ((m_rst >> e) & 1u) ? s.pos[e][d] : b.prev_targets[base + f];
      float x;
      if (d >= 7) { x = scale_to(av, lo, hi); x = ema * x + one_m_ema * pv; x = clamp_to(x, lo, hi); }
      else x = clamp_to(pv + gain * av, lo, hi);
      b.cur_targets[base + f] = x;
      b.prev_targets[base + f] = x;
      if (copy_actions) b.actions[base + f] = av;
    }
  }

  // ---------------- warp 1: random forces on the object body ----------------
  if (warp == 1) {
    if (do_force) {
      // Bernoulli test per env (lane = env), then (env, component) lanes apply it (:1401-1414)
      bool hit = false;
      if (valid) {
        const float prob = rst ? force_prob_from(p, R.u4(env, HA_KUKA_RND_FORCE_PROB).x) : f_prob;
        hit = R.u(env, HA_KUKA_RND_FORCE_TEST) < prob;
      }
      const unsigned m_hit = __ballot_sync(0xffffffffu, hit);
      const float decay = p.force_decay_pow, g0 = p.force_gain[0], g1 = p.force_gain[1];
#pragma unroll
      for (int i = 0; i < 3; ++i) {
        const int idx = lane + 32 * i, e = idx / 3, c = idx - 3 * e;
        if (e < nvalid) {
          float v = ((m_rst >> e) & 1u) ? 0.f : f_v[i];          // rb_forces[env_ids] = 0 (:1255)
          v *= decay;                                            // (:1402)
          if ((m_hit >> e) & 1u) v = R.nrm(env0 + e, HA_KUKA_RND_FORCE_NORMAL, c) * g0 * g1;   // (:1405-1410)
          b.rb_forces[frow + (size_t)e * fstride + c] = v;
        }
      }
    } else if (rst) {
      float* f = b.rb_forces + frow + (size_t)lane * fstride;
      f[0] = 0.f; f[1] = 0.f; f[2] = 0.f;
    }
  }

  // ---------------- warps 2,3: zero per-episode accumulators ---------------
  if (warp >= 2 && rst) {
#pragma unroll
    for (int k = warp - 2; k < HA_KUKA_N_EPISODE_TERMS; k += 2) b.rewards_episode[(size_t)k * n + env] = 0.0f;   // (:1341-1342)
  }

  // ---------------- warp 0: per-env scalar reset work (tiles with a reset only) ----------------
  if (warp == 0 && tile_resets) kuka_reset_apply(p, b, pf, env, rst, grst);
}

// advances the fast-mode RNG counter after an out-of-step reset (inside a step ha_kuka_post_step does it)
__global__ void kuka_rng_tick_kernel(HaKukaDynamic* dyn) { dyn->rng_step = dyn->rng_step + 1; }

}  // namespace ha
