// K_post for the in-hand reorientation family (ShadowHand, AllegroHand): one launch replaces
// post_physics_step (shadow_hand.py:710-715 / allegro_hand.py:627-632) = compute_observations
// (:437-584 / :406-504; observation layouts openai / full_no_vel / full / full_state and the asymmetric
// states_buf) + compute_hand_reward (:746-800 / :663-719) + the VecTask.step tail (vec_task.py:424-432).
//
// Mapping: CTA = 32-env tile, 128 threads.
//   load      cp.async into shared memory: 16-byte copies of the contiguous tiles (DOF state, DOF forces,
//             force sensors, actions), 4-byte copies of the 52-byte AoS rows (object root row, goal pose,
//             fingertip rigid-body rows);
//   describe  the observation row is a concatenation of scaled copies of those sources: every thread builds
//             the descriptors of its columns once (source offset, per-env stride, op) in shared memory, so
//             ONE code path serves all layouts of both hands and the asymmetric state row;
//   reward    warp 0, lane = env: quaternion difference (also an observation), compute_hand_reward, per-env
//             stores, statistics partials; the global consecutive_successes average is finished by the last CTA
//             from integer accumulators (order-independent);
//   assemble  warp per env row, lane = column (mod 32) with the column descriptors in registers: apply the column op
//             and the +-clipObservations clip and write 128-byte coalesced row segments straight to obs_buf.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"

namespace ha {

enum { kHandPostFull = 0, kHandPostObserveOnly = 1 };
constexpr int kHandThreads = 128;
constexpr int kHandTile = 32;
constexpr int kHandNPart = HA_HAND_N_PARTIAL_STATS;
constexpr int kHandFoldPerThread = HA_HAND_STAT_SLOTS / 128;      // the folding kernels run 128 threads
static_assert(HA_HAND_STAT_SLOTS % 128 == 0, "slot rows are folded by 128 threads");

// ---- deferred statistics (grid <= HA_HAND_STAT_SLOTS): the post kernel leaves one slot of integer partials per CTA in
// stat_accum[stat][slot] (plain stores, no atomics, no last-CTA tail) and raises tickets[1]; CTA s < kHandNPart of the next
// pre kernel (or of hand_stats_finalize_kernel) folds row s into stats[s].  The CTA of SUM_FINISHED_SUCCESSES also folds
// SUM_RESETS and applies the consecutive_successes average ONCE per post step (tickets[1] is its pending flag).
struct HandStatFold { long long v[kHandFoldPerThread], w[kHandFoldPerThread]; };

__device__ __forceinline__ void hand_fold_load(const long long* accum, const int stat, const int tid, HandStatFold& f) {
  const long long* row = accum + (size_t)stat * HA_HAND_STAT_SLOTS;
  const long long* row2 = accum + (size_t)HA_HAND_STAT_SUM_RESETS * HA_HAND_STAT_SLOTS;
#pragma unroll
  for (int i = 0; i < kHandFoldPerThread; ++i) {
    f.v[i] = row[tid + 128 * i];
    f.w[i] = (stat == HA_HAND_STAT_SUM_FINISHED_SUCCESSES) ? row2[tid + 128 * i] : 0;
  }
}
// red: 8 shared-memory entries (4 warps x {v, w})
__device__ __forceinline__ void hand_fold_warp(const HandStatFold& f, const int tid, long long* red) {
  long long v = 0, w = 0;
#pragma unroll
  for (int i = 0; i < kHandFoldPerThread; ++i) { v += f.v[i]; w += f.w[i]; }
  v = warp_sum_ll(v);
  w = warp_sum_ll(w);
  if ((tid & 31) == 0) { red[tid >> 5] = v; red[4 + (tid >> 5)] = w; }
}
// after a CTA barrier, one thread.  av_factor / n_envs are the only task constants of the tail.
__device__ __forceinline__ void hand_fold_finish(float* stats, float* consecutive_successes, unsigned* tickets, const int stat,
                                                 const int n_envs, const float av, const long long* red) {
  const long long raw = red[0] + red[1] + red[2] + red[3];
  const float fv = (stat == HA_HAND_STAT_SUM_REWARD) ? (float)((double)raw / 65536.0) : (float)raw;
  stats[stat] = fv;
  if (stat == HA_HAND_STAT_SUM_REWARD) stats[HA_HAND_STAT_MEAN_REWARD] = fv / (float)n_envs;
  if (stat == HA_HAND_STAT_SUM_FINISHED_SUCCESSES && tickets[1] != 0u) {
    const float num_resets = (float)(red[4] + red[5] + red[6] + red[7]);
    const float cons = consecutive_successes[0];
    const float upd = av * fv / num_resets + (1.0f - av) * cons;
    const float out = (num_resets > 0.0f) ? upd : cons;
    consecutive_successes[0] = out;
    stats[HA_HAND_STAT_CONSECUTIVE_SUCCESSES] = out;
    tickets[1] = 0u;
  }
}


struct HandStageOfs {   // float offsets into the stage (contiguous images first: they are filled by 16-byte copies)
  int dof, frc, sens, act, obj, goal, tip, qd, total;
};
__host__ __device__ inline HandStageOfs hand_stage_ofs(int D, int A, int T) {
  HandStageOfs o{};
  int k = 0;
  o.dof = k; k += kHandTile * 2 * D;
  o.frc = k; k += kHandTile * D;
  o.sens = k; k += kHandTile * 6 * T;
  o.act = k; k += kHandTile * A;
  o.obj = k; k += kHandTile * 13;
  o.goal = k; k += kHandTile * 7;
  o.tip = k; k += kHandTile * 13 * T;
  o.qd = k; k += kHandTile * 4;
  o.total = (k + 3) & ~3;
  return o;
}
// dynamic shared memory of the post kernel: stage + two descriptor tables
__host__ __device__ inline size_t hand_post_smem_bytes(int D, int A, int T) {
  return (size_t)hand_stage_ofs(D, A, T).total * 4 + 2 * HA_HAND_MAX_OBS * sizeof(HandCol);
}

// Descriptor of column j of observation layout `type` (shadow_hand.py:473-584, allegro_hand.py:437-504)
__device__ inline HandCol hand_col(const HaHandParams& p, const HandStageOfs& o, int type, int j) {
  const int D = p.n_dofs, A = p.n_actions, T = p.is_shadow ? p.n_tips : 0;
  int k = j;
  auto copy = [](int base, int stride) { return HandCol{base, stride, kOpCopy, 0.f, 0.f, 0.f}; };
  auto scaled = [](int base, int stride, float s) { return HandCol{base, stride, kOpScale, s, 0.f, 0.f}; };
  if (type == HA_HAND_OBS_OPENAI) {                    // fingertip positions, object position, relative rotation, actions
    if (k < 3 * T) { const int t = k / 3, c = k - 3 * t; return copy(o.tip + t * 13 + c, 13 * T); }
    k -= 3 * T;
    if (k < 3) return copy(o.obj + k, 13);
    k -= 3;
    if (k < 4) return copy(o.qd + k, 4);
    k -= 4;
    return copy(o.act + k, A);
  }
  if (k < D) return HandCol{o.dof + 2 * k, 2 * D, kOpUnscale, p.dof_lower[k], p.dof_upper[k], p.dof_range_rcp[k]};
  k -= D;
  if (type != HA_HAND_OBS_FULL_NO_VEL) {
    if (k < D) return scaled(o.dof + 2 * k + 1, 2 * D, p.vel_obs_scale);
    k -= D;
    if (type == HA_HAND_OBS_FULL_STATE) {
      if (k < D) return scaled(o.frc + k, D, p.force_torque_obs_scale);
      k -= D;
    }
    if (k < 13) return k < 10 ? copy(o.obj + k, 13) : scaled(o.obj + k, 13, p.vel_obs_scale);   // pose, linvel, 0.2 * angvel
    k -= 13;
  } else {
    if (k < 7) return copy(o.obj + k, 13);
    k -= 7;
  }
  if (k < 7) return copy(o.goal + k, 7);
  k -= 7;
  if (k < 4) return copy(o.qd + k, 4);
  k -= 4;
  if (type == HA_HAND_OBS_FULL_NO_VEL) {
    if (k < 3 * T) { const int t = k / 3, c = k - 3 * t; return copy(o.tip + t * 13 + c, 13 * T); }
    k -= 3 * T;
  } else {
    if (k < 13 * T) return copy(o.tip + k, 13 * T);
    k -= 13 * T;
    if (type == HA_HAND_OBS_FULL_STATE) {
      if (k < 6 * T) return scaled(o.sens + k, 6 * T, p.force_torque_obs_scale);
      k -= 6 * T;
    }
  }
  return copy(o.act + k, A);
}

__global__ void __launch_bounds__(kHandThreads)
hand_post_kernel(const __grid_constant__ HaHandParams p, const __grid_constant__ HaHandBuffers b, const int mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int D = p.n_dofs, A = p.n_actions, T = p.is_shadow ? p.n_tips : 0, Dp = p.dofs_per_env;
  const HandStageOfs o = hand_stage_ofs(D, A, T);
  float* stage = reinterpret_cast<float*>(smem_raw);
  HandCol* cols_obs = reinterpret_cast<HandCol*>(stage + o.total);
  HandCol* cols_st = cols_obs + HA_HAND_MAX_OBS;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.n_envs;
  const int env0 = blockIdx.x * kHandTile;
  const int nvalid = min(kHandTile, n - env0), last = nvalid - 1;
  const bool has_states = p.states_dim > 0;
  const bool need_frc = (p.obs_type == HA_HAND_OBS_FULL_STATE) || has_states;
  const bool need_dof = (p.obs_type != HA_HAND_OBS_OPENAI) || has_states;

  // ---------------- load ----------------
  for (int j = tid; j < kHandTile * 13; j += kHandThreads) {          // object root rows
    const int e = j / 13, c = j - 13 * e, env = env0 + min(e, last);
    const size_t row = p.root_regular ? (size_t)env * p.actors_per_env + p.object_slot : (size_t)__ldg(b.object_actor + env);
    cp_async4(stage + o.obj + j, b.root_state + row * 13 + c);
  }
  for (int j = tid; j < kHandTile * 7; j += kHandThreads) {           // goal pose rows
    const int e = j / 7, c = j - 7 * e;
    cp_async4(stage + o.goal + j, b.goal_states + (size_t)(env0 + min(e, last)) * 13 + c);
  }
  if (T > 0) {                                                        // fingertip rigid-body rows (all 13 floats)
    const size_t rbs = (size_t)p.n_bodies * 13;
    for (int j = tid; j < kHandTile * 13 * T; j += kHandThreads) {
      const int e = j / (13 * T), r = j - 13 * T * e, t = r / 13, c = r - 13 * t;
      cp_async4(stage + o.tip + j, b.rigid_body_state + (size_t)(env0 + min(e, last)) * rbs + (size_t)p.tip_body[t] * 13 + c);
    }
  }
  {
    // contiguous tiles; a full tile starts 16-byte aligned (env0 % 32 == 0) when the row widths keep it so
    auto tile = [&](float* dst, const float* src, int per_env) {
      const float* s0 = src + (size_t)env0 * per_env;
      const int count = nvalid * per_env;
      if (nvalid == kHandTile && ((per_env * kHandTile) & 3) == 0 && ((((size_t)env0 * per_env) & 3) == 0)) {
        for (int q = tid; q < (count >> 2); q += kHandThreads) cp_async16(reinterpret_cast<float4*>(dst) + q, reinterpret_cast<const float4*>(s0) + q);
      } else {
        for (int j = tid; j < count; j += kHandThreads) cp_async4(dst + j, s0 + j);
      }
    };
    if (need_dof) {
      if (Dp == D) tile(stage + o.dof, b.dof_state, 2 * D);
      else for (int j = tid; j < nvalid * 2 * D; j += kHandThreads) {   // env rows are wider than the hand's DOFs
        const int e = j / (2 * D), r = j - 2 * D * e;
        cp_async4(stage + o.dof + j, b.dof_state + (size_t)(env0 + e) * 2 * Dp + r);
      }
    }
    if (need_frc) {
      tile(stage + o.frc, b.dof_force, D);
      if (T > 0) tile(stage + o.sens, b.force_sensors, 6 * T);
    }
    tile(stage + o.act, b.actions, A);
  }
  cp_async_commit();
  // column descriptors of the two rows
  for (int j = tid; j < p.obs_dim; j += kHandThreads) cols_obs[j] = hand_col(p, o, p.obs_type, j);
  if (has_states)
    for (int j = tid; j < p.states_dim; j += kHandThreads) cols_st[j] = hand_col(p, o, HA_HAND_OBS_FULL_STATE, j);
  // per-env state of the reward warp (lane = env)
  const bool valid = lane < nvalid;
  const int env = env0 + min(lane, last);
  long long r_reset = 0, r_goal = 0, r_progress = 0, r_randomize = 0;
  float r_succ = 0.f;
  if (warp == 0 && mode == kHandPostFull) {
    r_reset = b.reset_buf[env]; r_goal = b.reset_goal_buf[env]; r_progress = b.progress_buf[env];
    r_succ = b.successes[env];
    if (b.randomize_buf) r_randomize = b.randomize_buf[env];
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---------------- reward (warp 0, lane = env) ----------------
  long long a_rew = 0;
  int a_n = 0, a_resets = 0, a_goal = 0, a_timeouts = 0;
  long long a_fin = 0, a_succ = 0;
  if (warp == 0) {
    const float* ob = stage + o.obj + lane * 13;
    const float* gl = stage + o.goal + lane * 7;
    const f4 qd = quat_mul(f4{ob[3], ob[4], ob[5], ob[6]}, quat_conjugate(f4{gl[3], gl[4], gl[5], gl[6]}));
    float* sq = stage + o.qd + lane * 4;
    sq[0] = qd.x; sq[1] = qd.y; sq[2] = qd.z; sq[3] = qd.w;
    if (mode == kHandPostFull) {
      const long long progress_in = r_progress + 1;                      // progress_buf += 1 (:711)
      const float goal_dist = norm3(f3{ob[0] - gl[0], ob[1] - gl[1], ob[2] - gl[2]});
      const float tol = p.ignore_z_rot ? 2.0f * p.success_tolerance : p.success_tolerance;
      const float rot_dist = 2.0f * asinf(fminf(norm3(f3{qd.x, qd.y, qd.z}), 1.0f));
      const float dist_rew = goal_dist * p.dist_reward_scale;
      const float rot_rew = 1.0f / (fabsf(rot_dist) + p.rot_eps) * p.rot_reward_scale;
      float ap = 0.f;
      const float* ac = stage + o.act + lane * A;
      for (int a = 0; a < A; ++a) ap += ac[a] * ac[a];
      float reward = dist_rew + rot_rew + ap * p.action_penalty_scale;
      const bool hit = fabsf(rot_dist) <= tol;
      const long long goal_resets = hit ? 1 : r_goal;
      const float succ = r_succ + (float)goal_resets;
      reward = (goal_resets == 1) ? reward + p.reach_goal_bonus : reward;
      const bool fell = goal_dist >= p.fall_dist;
      reward = fell ? reward + p.fall_penalty : reward;
      long long resets = fell ? 1 : r_reset;
      long long prog = progress_in;
      if (p.max_consecutive_successes > 0) {
        prog = hit ? 0 : prog;
        resets = (succ >= (float)p.max_consecutive_successes) ? 1 : resets;
      }
      const bool timed_out = prog >= (long long)(p.max_episode_length - 1);
      resets = timed_out ? 1 : resets;
      if (p.max_consecutive_successes > 0) reward = timed_out ? reward + 0.5f * p.fall_penalty : reward;
      const bool timeout_flag = timed_out && (resets != 0);              // vec_task.py:424
      if (valid) {
        b.rew_buf[env] = reward;
        b.reset_buf[env] = resets;
        b.reset_goal_buf[env] = goal_resets;
        b.progress_buf[env] = prog;
        b.successes[env] = succ;
        if (b.randomize_buf) b.randomize_buf[env] = r_randomize + 1;
        b.timeout_buf[env] = timeout_flag ? 1 : 0;
        a_rew = __float2ll_rn(reward * 65536.0f);
        a_n = 1; a_resets = (int)resets; a_goal = goal_resets != 0; a_timeouts = timeout_flag;     // num_resets = sum(resets)
        a_fin = (long long)__float2ll_rn(succ * (float)resets);          // successes * resets.float()
        a_succ = (long long)__float2ll_rn(succ);
      }
    }
  }
  __syncthreads();                                                       // quat_diff is in the stage

  // ---------------- assemble ----------------
  constexpr int kMaxPerLane = (HA_HAND_MAX_OBS + 31) / 32;
  assemble_rows<kMaxPerLane>(stage, cols_obs, b.obs_buf + (size_t)env0 * p.obs_dim, p.obs_dim, nvalid, p.clip_obs, lane, warp, kHandThreads / 32);
  if (has_states)
    assemble_rows<kMaxPerLane>(stage, cols_st, b.states_buf + (size_t)env0 * p.states_dim, p.states_dim, nvalid, p.clip_obs, lane, warp, kHandThreads / 32);

  // ---------------- statistics + the global consecutive_successes average (warp 0) ----------------
  if (warp != 0 || mode != kHandPostFull) return;
  const unsigned FULL = 0xffffffffu;
  const long long s_rew = warp_sum_ll(a_rew), s_fin = warp_sum_ll(a_fin), s_succ = warp_sum_ll(a_succ);
  const int s_n = warp_sum_i(a_n), s_rs = warp_sum_i(a_resets), s_gl = warp_sum_i(a_goal), s_to = warp_sum_i(a_timeouts);
  long long* accum = reinterpret_cast<long long*>(b.stat_accum);
  long long val = 0;
  // selects, not a switch: a switch on the lane index is a divergent jump table (one serial pass per case)
  val = lane == HA_HAND_STAT_N ? (long long)(s_n) : val;
  val = lane == HA_HAND_STAT_SUM_REWARD ? (long long)(s_rew) : val;
  val = lane == HA_HAND_STAT_SUM_RESETS ? (long long)(s_rs) : val;
  val = lane == HA_HAND_STAT_SUM_FINISHED_SUCCESSES ? (long long)(s_fin) : val;
  val = lane == HA_HAND_STAT_SUM_GOAL_RESETS ? (long long)(s_gl) : val;
  val = lane == HA_HAND_STAT_SUM_SUCCESSES ? (long long)(s_succ) : val;
  val = lane == HA_HAND_STAT_SUM_TIMEOUTS ? (long long)(s_to) : val;
  if (gridDim.x <= HA_HAND_STAT_SLOTS) {                                        // deferred: see hand_fold_*
    if (lane < kHandNPart) accum[(size_t)lane * HA_HAND_STAT_SLOTS + blockIdx.x] = val;
    if (blockIdx.x == 0 && lane == 0) {
      b.tickets[1] = 1u;
      b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
      b.dynamic->rng_step = b.dynamic->rng_step + 1;
    }
    return;
  }
  long long old = 0;
  if (lane < HA_HAND_N_PARTIAL_STATS) old = (long long)atomicAdd(reinterpret_cast<unsigned long long*>(accum + lane), (unsigned long long)val);
  const unsigned dep = __any_sync(FULL, old == 0x7fffffffffffffffLL) ? 1u : 0u;   // ticket depends on every returned atomic
  int is_last = 0;
  if (lane == 0) is_last = (atomicAdd(b.tickets + 0, 1u + dep) == gridDim.x - 1);
  is_last = __shfl_sync(FULL, is_last, 0);
  if (!is_last) return;
  __threadfence();
  float fv = 0.f;
  if (lane < HA_HAND_N_PARTIAL_STATS) {
    const long long raw = (long long)atomicExch(reinterpret_cast<unsigned long long*>(accum + lane), 0ULL);   // read and re-arm
    fv = (lane == HA_HAND_STAT_SUM_REWARD) ? (float)((double)raw / 65536.0) : (float)raw;
    b.stats[lane] = fv;
  }
  const float cnt = __shfl_sync(FULL, fv, HA_HAND_STAT_N), sum_rew = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_REWARD);
  const float num_resets = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_RESETS), finished = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_FINISHED_SUCCESSES);
  if (lane == 0) {
    // cons = where(num_resets > 0, av * finished / num_resets + (1 - av) * cons, cons)   (:795-798 / :714-717)
    const float cons = b.consecutive_successes[0];
    const float av = p.av_factor;
    const float upd = av * finished / num_resets + (1.0f - av) * cons;
    const float out = (num_resets > 0.0f) ? upd : cons;
    b.consecutive_successes[0] = out;
    b.stats[HA_HAND_STAT_MEAN_REWARD] = sum_rew / cnt;
    b.stats[HA_HAND_STAT_CONSECUTIVE_SUCCESSES] = out;
    b.tickets[0] = 0;
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    b.dynamic->rng_step = b.dynamic->rng_step + 1;
  }
}

// the fold as a launch of its own (ha_hand_stats_finalize): grid kHandNPart x 128 threads
__global__ void __launch_bounds__(128) hand_stats_finalize_kernel(const __grid_constant__ HaHandBuffers b, const int n_envs, const float av) {
  __shared__ long long red[8];
  HandStatFold f;
  hand_fold_load(reinterpret_cast<const long long*>(b.stat_accum), blockIdx.x, threadIdx.x, f);
  hand_fold_warp(f, threadIdx.x, red);
  __syncthreads();
  if (threadIdx.x == 0) hand_fold_finish(b.stats, b.consecutive_successes, b.tickets, blockIdx.x, n_envs, av, red);
}

}  // namespace ha
