// K_post for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1426-1443 (post_physics_step) = compute_observations
// (:991-1089) + compute_full_state (:1091-1172) + compute_kuka_reward (:854-930)
// + _compute_resets (:841-849) + clamp_obs (:1174-1176), plus the VecTask.step
// tail (vec_task.py:424-432: timeout_buf, obs clip).
//
// Mapping (memory-bound; the limiter after DRAM is instruction issue, so the
// design minimises instructions per env and never repeats per-env work):
//   one WARP owns a tile of 32 consecutive envs, lane = env, CTA = one warp, no
//   CTA barrier; 255 registers per thread are available at this occupancy
//   (7 warps / SM cover 32768 envs in one wave), so loads are batched deep.
//   batch A  AoS rows (13-float rigid-body / root rows at 52-byte stride, only
//            4-byte aligned): 16 lanes per 13-float row, 8 lanes per 7-float
//            row, i.e. 2..4 rows per request; keypoint table, scales and the
//            per-env scalars.  All issued, then scattered to padded smem.
//   batch B  DOF state (contiguous, 128-bit) and the 12 episode accumulators:
//            issued BEFORE the geometry runs, consumed after it, so the second
//            batch's latency hides behind the quaternion math.
//   compute  lane = env: palm centre, 4 fingertips, K keypoints, reward and
//            reset logic in registers; observation row assembled in smem.
//   store    the 32 x obs_dim tile leaves as one contiguous run of 128-bit
//            stores; per-env outputs are coalesced 128/256-byte stores.
// Episode statistics: warp shuffle -> 64-bit fixed-point atomics, so sums are
// order-independent (bit-reproducible); the last warp converts them.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"

namespace ha {

struct KukaLayout {            // obs slot offsets, allegro_kuka_base.py:1091-1172
  int dof_pos, dof_vel, palm_pos, palm_rot_vel, obj_rot_vel, tip_rel_palm, kp_rel_palm, kp_rel_goal,
      scales, closest_kp, closest_tip, lifted, progress, successes, reward, dim;
};
__host__ __device__ constexpr KukaLayout kuka_layout(int n_dofs, int n_tips, int K) {
  KukaLayout l{};
  int o = 0;
  l.dof_pos = o; o += n_dofs;
  l.dof_vel = o; o += n_dofs;
  l.palm_pos = o; o += 3;
  l.palm_rot_vel = o; o += 10;
  l.obj_rot_vel = o; o += 10;
  l.tip_rel_palm = o; o += 3 * n_tips;
  l.kp_rel_palm = o; o += 3 * K;
  l.kp_rel_goal = o; o += 3 * K;
  l.scales = o; o += 3;
  l.closest_kp = o; o += 1;
  l.closest_tip = o; o += n_tips;
  l.lifted = o; o += 1;
  l.progress = o; o += 1;
  l.successes = o; o += 1;
  l.reward = o; o += 1;
  l.dim = o;
  return l;
}

enum { kPostFull = 0, kPostObserveOnly = 1 };

constexpr int kD1 = 23;              // single arm DOFs
constexpr int kT1 = 4;               // fingertips
constexpr int kTile = 32;            // envs per warp (lane = env)
constexpr int kPostThreads = 32;     // one warp per CTA
constexpr int kNPart = HA_KUKA_N_PARTIAL_STATS;

__host__ __device__ constexpr bool stat_is_max(int i) {
  return i == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE || i == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE;
}
// fixed-point scale of the 64-bit statistic accumulators
__host__ __device__ constexpr float stat_scale(int i) {
  return (i == HA_KUKA_STAT_SUM_REWARD) ? 65536.0f : (i == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE) ? 1048576.0f : 1.0f;
}
// order-preserving float <-> int64 key (for atomicMax on signed integers)
__device__ __forceinline__ long long float_key(float f) {
  const int b = __float_as_int(f);
  return (long long)(b >= 0 ? b : (b ^ 0x7fffffff));
}
__device__ __forceinline__ float key_float(long long k) {
  const int b = (int)k;
  return __int_as_float(b >= 0 ? b : (b ^ 0x7fffffff));
}
constexpr long long kStatMaxInit = -0x7fffffffLL - 1;   // below every float key

template <int K>
struct KukaPostSmem {                 // one warp; row strides are odd so that lane = env reads are conflict-free
  static constexpr int kObs = kuka_layout(kD1, kT1, K).dim;
  static constexpr int kTipW = kT1 * 7 + 1;
  static constexpr int kKpW = (3 * K) | 1;
  static constexpr int kStageA = kTile * (7 + kTipW + 7 + 7 + kKpW);
  static constexpr int kStageB = kTile * 2 * kD1;          // DOF tile; reuses the batch-A staging area
  static constexpr int kStage = ((kStageA > kStageB ? kStageA : kStageB) + 3) & ~3;
  float obs[kTile * kObs];            // row-major tile, leaves as one contiguous run
  float stage[kStage];
  __device__ __forceinline__ float* palm() { return stage; }                       // [e * 7 + c]   pos3 quat4
  __device__ __forceinline__ float* tip() { return stage + kTile * 7; }            // [e * kTipW + t * 7 + c]
  __device__ __forceinline__ float* obj() { return tip() + kTile * kTipW; }        // [e * 7 + c]
  __device__ __forceinline__ float* goal() { return obj() + kTile * 7; }           // [e * 7 + c]
  __device__ __forceinline__ float* kp() { return goal() + kTile * 7; }            // [e * kKpW + 3 k + c]
};

struct KukaPostEnvStats {   // per-env results that feed the statistics (idle lanes: valid == false)
  float reward = 0.f, prev_ep = 0.f, tobj = 0.f;
  bool valid = false, resets = false, success = false, timeout = false, lifted = false;
};

// Episode statistics of one warp's tile -> global fixed-point accumulators; the last warp of the grid
// converts them to floats.  Runs BEFORE the warp's bulk stores so that the fence that orders the
// atomics ahead of the ticket does not have to wait for 15 KB of observation stores to drain.
__device__ __forceinline__ void kuka_post_stats(const HaKukaBuffers& b, const KukaPostEnvStats& st, const int lane) {
  const unsigned FULL = 0xffffffffu;
  // 0/1 flags by ballot, the rest by a warp butterfly; lane i < kNPart then owns statistic i
  const bool c = st.valid;
  const unsigned m_n = __ballot_sync(FULL, c), m_rs = __ballot_sync(FULL, c && st.resets);
  const unsigned m_su = __ballot_sync(FULL, c && st.success), m_to = __ballot_sync(FULL, c && st.timeout);
  const unsigned m_li = __ballot_sync(FULL, c && st.lifted);
  float v_rew = c ? st.reward : 0.f, v_pe = c ? st.prev_ep : 0.f, v_to = c ? st.tobj : 0.f;
  float v_mx = c ? st.tobj : -3.0e38f, v_mn = c ? -st.tobj : -3.0e38f;
#pragma unroll
  for (int ofs = 16; ofs > 0; ofs >>= 1) {
    v_rew += __shfl_xor_sync(FULL, v_rew, ofs);
    v_pe += __shfl_xor_sync(FULL, v_pe, ofs);
    v_to += __shfl_xor_sync(FULL, v_to, ofs);
    v_mx = fmaxf(v_mx, __shfl_xor_sync(FULL, v_mx, ofs));
    v_mn = fmaxf(v_mn, __shfl_xor_sync(FULL, v_mn, ofs));
  }
  long long* accum = reinterpret_cast<long long*>(b.stat_accum);
  {
    float val = 0.f;
    switch (lane) {
      case HA_KUKA_STAT_N: val = (float)__popc(m_n); break;
      case HA_KUKA_STAT_SUM_REWARD: val = v_rew; break;
      case HA_KUKA_STAT_SUM_RESETS: val = (float)__popc(m_rs); break;
      case HA_KUKA_STAT_SUM_SUCCESS_EVENTS: val = (float)__popc(m_su); break;
      case HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES: val = v_pe; break;
      case HA_KUKA_STAT_SUM_TRUE_OBJECTIVE: val = v_to; break;
      case HA_KUKA_STAT_MAX_TRUE_OBJECTIVE: val = v_mx; break;
      case HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE: val = v_mn; break;
      case HA_KUKA_STAT_SUM_TIMEOUTS: val = (float)__popc(m_to); break;
      case HA_KUKA_STAT_SUM_LIFTED: val = (float)__popc(m_li); break;
      default: break;
    }
    if (lane < kNPart) {
      if (stat_is_max(lane)) atomicMax(accum + lane, float_key(val));
      else atomicAdd(reinterpret_cast<unsigned long long*>(accum + lane), (unsigned long long)__float2ll_rn(val * stat_scale(lane)));
    }
  }
  __syncwarp();
  int is_last = 0;
  if (lane == 0) {
    __threadfence();                                  // this warp's atomics are ordered before its ticket
    is_last = (atomicAdd(b.tickets + 0, 1u) == gridDim.x - 1);
  }
  is_last = __shfl_sync(FULL, is_last, 0);
  if (!is_last) return;
  // ---------------------- last warp: final statistics ----------------------
  __threadfence();
  float val = 0.f;
  if (lane < kNPart) {
    const bool is_max = stat_is_max(lane);
    const long long raw = (long long)atomicExch(reinterpret_cast<unsigned long long*>(accum + lane),
                                                (unsigned long long)(is_max ? kStatMaxInit : 0LL));   // read and re-arm
    val = is_max ? key_float(raw) : (float)((double)raw / (double)stat_scale(lane));
    b.stats[lane] = val;
  }
  // means / min for extras (allegro_kuka_base.py:903-910), so the host needs no extra launches
  const float cnt = __shfl_sync(FULL, val, HA_KUKA_STAT_N);
  const float s_rew = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_REWARD), s_pe = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES);
  const float s_to = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_TRUE_OBJECTIVE), s_mn = __shfl_sync(FULL, val, HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE);
  if (lane == 0) {
    const float inv_n = 1.0f / cnt;
    b.stats[HA_KUKA_STAT_MEAN_REWARD] = s_rew * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES] = s_pe * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE] = s_to * inv_n;
    b.stats[HA_KUKA_STAT_MIN_TRUE_OBJECTIVE] = -s_mn;
    b.tickets[0] = 0;                    // re-arm for the next launch (graph replay safe)
    // the reset/index counters of the pre-step are consumed by now: clear for the next pre-step
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    // fast-mode RNG: one counter tick per step
    b.dynamic->rng_step = b.dynamic->rng_step + 1;
  }
}

// One warp, one tile of up to 32 envs, lane = env.  kFull: all 32 envs exist (no index clamps).
template <int K, bool kFull>
__device__ __forceinline__ void kuka_post_tile(const HaKukaParams& p, const HaKukaBuffers& b, KukaPostSmem<K>& s, const int mode,
                                               const int env0, const int nvalid_rt, const int lane) {
  using S = KukaPostSmem<K>;
  constexpr int OBS = S::kObs, TIPW = S::kTipW, KPW = S::kKpW;
  constexpr KukaLayout L = kuka_layout(kD1, kT1, K);
  const int n = p.n_envs;
  const int nvalid = kFull ? kTile : nvalid_rt;
  const bool valid = kFull ? true : (lane < nvalid);
  const int last = nvalid - 1;
  auto cl = [&](int e) { return kFull ? e : min(e, last); };   // idle lanes of a ragged tile re-read a valid env
  const int env = env0 + cl(lane);                             // stores are predicated on `valid`

  // ------------------------------ batch A: issue ---------------------------------
  // AoS rows.  16 lanes per 13-float row (2 rows / request), 8 lanes per 7-float row (4 rows / request).
  const int h16 = lane >> 4, c16 = lane & 15, c16c = min(c16, 12);
  const int r8 = lane >> 3, c8 = lane & 7, c8c = min(c8, 6);
  const size_t rb_env_stride = (size_t)p.n_bodies * 13;
  const float* rb0 = b.rigid_body_state + (size_t)env0 * rb_env_stride;
  float palm_v[16], obj_v[16], tip_v[32], goal_v[8];
  {
    const float* src = rb0 + (size_t)p.palm_body[0] * 13 + c16c;
#pragma unroll
    for (int i = 0; i < 16; ++i) palm_v[i] = ldg(src + (size_t)cl(2 * i + h16) * rb_env_stride);
  }
  if (p.root_regular) {
    const float* src = b.root_state + ((size_t)env0 * p.actors_per_env + p.object_slot) * 13 + c16c;
    const size_t stride = (size_t)p.actors_per_env * 13;
#pragma unroll
    for (int i = 0; i < 16; ++i) obj_v[i] = src[(size_t)cl(2 * i + h16) * stride];
  } else {
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const int row = __ldg(b.object_actor + env0 + cl(2 * i + h16));
      obj_v[i] = b.root_state[(size_t)row * 13 + c16c];
    }
  }
  {
    const float* src = rb0 + (size_t)p.tip_body[r8] * 13 + c8c;
#pragma unroll
    for (int i = 0; i < 32; ++i) tip_v[i] = ldg(src + (size_t)cl(i) * rb_env_stride);
  }
  {
    const float* src = b.goal_states + (size_t)env0 * 13 + c8c;
#pragma unroll
    for (int i = 0; i < 8; ++i) goal_v[i] = src[(size_t)cl(4 * i + r8) * 13];
  }
  // keypoint table [nvalid, K, 3] and scales [nvalid, 3]: contiguous
  constexpr int kKpIters = 3 * K;                     // kTile * 3K floats / 32 lanes
  float kp_v[kKpIters];
  float4 kp_v4[3];
  if constexpr (K == 4) {                             // 12 floats per env: whole tile is float4-aligned, nvalid * 12 % 4 == 0
    const float4* src = reinterpret_cast<const float4*>(b.object_keypoint_offsets + (size_t)env0 * 12);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int q = lane + 32 * i;
      kp_v4[i] = (kFull || q < nvalid * 3) ? __ldg(src + q) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  } else {
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = lane + 32 * i;
      kp_v[i] = (kFull || idx < nvalid * 3 * K) ? ldg(b.object_keypoint_offsets + (size_t)env0 * 3 * K + idx) : 0.f;
    }
  }
  float scale_v[3];
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const int idx = lane + 32 * i;
    scale_v[i] = (kFull || idx < nvalid * 3) ? ldg(b.object_scales + (size_t)env0 * 3 + idx) : 0.f;
  }
  // per-env scalars, lane = env: every request is one coalesced 128 / 256-byte run
  const long long r_progress = b.progress_buf[env];
  const long long r_reset = b.reset_buf[env];
  const float r_succ = b.successes[env];
  const float r_ckp = b.closest_keypoint_max_dist[env];
  const float r_furthest = b.furthest_hand_dist[env];
  const int r_near = b.near_goal_steps[env];
  const unsigned char r_lifted = b.lifted_object[env];
  const float r_initz = ldg(b.object_init_state + (size_t)env * 13 + 2);
  const float r_prev_ep = b.prev_episode_successes[env];
  const float4 r_ctip = *reinterpret_cast<const float4*>(b.closest_fingertip_dist + (size_t)env * 4);
  const float4 r_coef = __ldg(reinterpret_cast<const float4*>(b.finger_rew_coeffs + (size_t)env * 4));
  float r_old_rew_obs = 0.f;
  if (mode == kPostObserveOnly) r_old_rew_obs = b.obs_buf[(size_t)env * OBS + L.reward];
  const float keypoint_tol = b.dynamic->keypoint_success_tolerance;
  const float tobj_scale = b.dynamic->true_objective_scale, tobj_offset = b.dynamic->true_objective_offset;

  // ------------------------------ batch A: scatter to smem ---------------------------------
  float* s_palm = s.palm(); float* s_tip = s.tip(); float* s_obj = s.obj(); float* s_goal = s.goal(); float* s_kp = s.kp();
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const int e = 2 * i + h16;
    if (c16 < 7) { s_palm[e * 7 + c16] = palm_v[i]; s_obj[e * 7 + c16] = obj_v[i]; }
    if (c16 >= 3 && c16 < 13) {                       // palm rot + vel (:1112-1117), object rot + vel (:1120)
      s.obs[e * OBS + L.palm_rot_vel + c16 - 3] = palm_v[i];
      s.obs[e * OBS + L.obj_rot_vel + c16 - 3] = obj_v[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 32; ++i)
    if (c8 < 7) s_tip[i * TIPW + r8 * 7 + c8] = tip_v[i];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    if (c8 < 7) s_goal[(4 * i + r8) * 7 + c8] = goal_v[i];
  if constexpr (K == 4) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {                     // float4 q covers floats 4q..4q+3 of env q / 3, keypoint (q % 3) * 4 / 3 ...
      const int q = lane + 32 * i, e = q / 3, o4 = 4 * (q - 3 * e);
      float* d = s_kp + e * KPW + o4;
      d[0] = kp_v4[i].x; d[1] = kp_v4[i].y; d[2] = kp_v4[i].z; d[3] = kp_v4[i].w;
    }
  } else {
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = lane + 32 * i, e = idx / (3 * K);
      s_kp[e * KPW + (idx - e * 3 * K)] = kp_v[i];
    }
  }
#pragma unroll
  for (int i = 0; i < 3; ++i) {                       // object scales (:1143)
    const int idx = lane + 32 * i, e = idx / 3;
    s.obs[e * OBS + L.scales + (idx - 3 * e)] = scale_v[i];
  }
  __syncwarp();

  // ------------------------------ batch B: issue (consumed after the geometry) ---------------------------------
  const float* dof_src = b.dof_state + (size_t)env0 * (2 * kD1);
  const int dof_total = nvalid * 2 * kD1, dof_nvec = dof_total >> 2;      // tile base is 16-byte aligned (env0 % 32 == 0)
  constexpr int kDofIters = (kTile * 2 * kD1 / 4 + 31) / 32;              // 12
  float4 dof_v[kDofIters];
#pragma unroll
  for (int i = 0; i < kDofIters; ++i) {
    const int q = lane + 32 * i;
    dof_v[i] = (q < dof_nvec) ? __ldg(reinterpret_cast<const float4*>(dof_src) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  float dof_tail = 0.f;                               // odd nvalid leaves 2 floats after the last float4
  if (!kFull && (dof_nvec << 2) + lane < dof_total) dof_tail = ldg(dof_src + (dof_nvec << 2) + lane);
  float r_acc[HA_KUKA_N_EPISODE_TERMS];
  long long r_randomize = 0;
  if (mode == kPostFull) {
#pragma unroll
    for (int k = 0; k < HA_KUKA_N_EPISODE_TERMS; ++k) r_acc[k] = b.rewards_episode[(size_t)k * n + env];
    if (b.randomize_buf) r_randomize = b.randomize_buf[env];
  }

  // ------------------------------ geometry, lane = env ---------------------------------
  float* o = s.obs + lane * OBS;
  const float* pr = s_palm + lane * 7;
  const float* ob = s_obj + lane * 7;
  // palm_center_pos = palm_pos + quat_rotate(palm_rot, palm_offset)   (:1013-1017)
  const f3 pc = f3{pr[0], pr[1], pr[2]} + quat_rotate(f4{pr[3], pr[4], pr[5], pr[6]}, f3{p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]});
  const f3 obj_pos = {ob[0], ob[1], ob[2]};
  const f4 obj_rot = {ob[3], ob[4], ob[5], ob[6]};
  const float obj_z = ob[2];
  o[L.palm_pos + 0] = pc.x; o[L.palm_pos + 1] = pc.y; o[L.palm_pos + 2] = pc.z;
  float cur[kT1];
#pragma unroll
  for (int t = 0; t < kT1; ++t) {   // fingertips (:1033-1052)
    const float* tr = s_tip + lane * TIPW + t * 7;
    const f3 to = f3{tr[0], tr[1], tr[2]} +
                  quat_rotate(f4{tr[3], tr[4], tr[5], tr[6]}, f3{p.tip_offset[t][0], p.tip_offset[t][1], p.tip_offset[t][2]});
    const f3 rel_obj = to - obj_pos, rel_palm = to - pc;
    cur[t] = norm3(rel_obj);
    o[L.tip_rel_palm + 3 * t + 0] = rel_palm.x;
    o[L.tip_rel_palm + 3 * t + 1] = rel_palm.y;
    o[L.tip_rel_palm + 3 * t + 2] = rel_palm.z;
  }
  float kp_max = 0.0f;
  {
    const float* gl = s_goal + lane * 7;
    const f3 goal_pos = {gl[0], gl[1], gl[2]};
    const f4 goal_rot = {gl[3], gl[4], gl[5], gl[6]};
#pragma unroll
    for (int k = 0; k < K; ++k) {   // keypoints (:1057-1073)
      const float* kp = s_kp + lane * KPW + 3 * k;
      const f3 off = {kp[0], kp[1], kp[2]};
      const f3 okp = obj_pos + quat_rotate(obj_rot, off);
      const f3 gkp = goal_pos + quat_rotate(goal_rot, off);
      const f3 rel_goal = okp - gkp, rel_palm = okp - pc;
      const float dist = norm3(rel_goal);
      kp_max = (k == 0) ? dist : fmaxf(kp_max, dist);
      o[L.kp_rel_palm + 3 * k + 0] = rel_palm.x;
      o[L.kp_rel_palm + 3 * k + 1] = rel_palm.y;
      o[L.kp_rel_palm + 3 * k + 2] = rel_palm.z;
      o[L.kp_rel_goal + 3 * k + 0] = rel_goal.x;
      o[L.kp_rel_goal + 3 * k + 1] = rel_goal.y;
      o[L.kp_rel_goal + 3 * k + 2] = rel_goal.z;
    }
  }
  float ctip[4] = {r_ctip.x, r_ctip.y, r_ctip.z, r_ctip.w};
  const float coef[4] = {r_coef.x, r_coef.y, r_coef.z, r_coef.w};
  // lazy initialisation of the per-episode closest / furthest distances (:1044-1049, :1077-1079)
#pragma unroll
  for (int i = 0; i < 4; ++i) ctip[i] = (ctip[i] < 0.0f) ? cur[i] : ctip[i];
  float furthest = (r_furthest < 0.0f) ? cur[0] : r_furthest;
  float ckp = (r_ckp < 0.0f) ? kp_max : r_ckp;
  const long long progress = r_progress + (mode == kPostFull ? 1 : 0);   // progress_buf += 1 (:1429)
  // observation slots that see pre-update state (ordering: SURVEY 8a)
  o[L.closest_kp] = ckp;
#pragma unroll
  for (int i = 0; i < 4; ++i) o[L.closest_tip + i] = ctip[i];
  o[L.lifted] = r_lifted ? 1.0f : 0.0f;
  o[L.progress] = logf((float)progress / 10.0f + 1.0f);
  o[L.successes] = logf(r_succ + 1.0f);

  // ------------------------------ batch B: DOF tile through smem ---------------------------------
  __syncwarp();                                       // every lane is done with the batch-A staging area
  {
    float4* sd = reinterpret_cast<float4*>(s.stage);
#pragma unroll
    for (int i = 0; i < kDofIters; ++i) {
      const int q = lane + 32 * i;
      if (q < kTile * 2 * kD1 / 4) sd[q] = dof_v[i];
    }
    if (!kFull && (dof_nvec << 2) + lane < dof_total) s.stage[(dof_nvec << 2) + lane] = dof_tail;
  }
  __syncwarp();
  float sa = 0.f, sh = 0.f;                           // _action_penalties sums (:830-839)
  {
    const float2* dl = reinterpret_cast<const float2*>(s.stage) + lane * kD1;   // (pos, vel) pairs of this env
#pragma unroll
    for (int d = 0; d < kD1; ++d) {
      const float2 pv = dl[d];
      // unscale(dof_pos) (:1095-1100), raw dof_vel (:1104); limits come straight from the constant bank
      o[L.dof_pos + d] = unscale_from(pv.x, p.dof_lower[d], p.dof_upper[d]);
      o[L.dof_vel + d] = pv.y;
      if (d < 7) sa += fabsf(pv.y); else sh += fabsf(pv.y);
    }
  }

  if (mode == kPostObserveOnly) {
    o[L.reward] = r_old_rew_obs;
    if (valid) {
      b.closest_keypoint_max_dist[env] = ckp;
      b.furthest_hand_dist[env] = furthest;
      *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
    }
  } else {
    // ------------------------ reward / reset logic ---------------------------
    // _lifting_reward (:788-812)
    const float z_lift = 0.05f + obj_z - r_initz;
    float lifting_rew = clipf(z_lift, 0.0f, 0.5f);
    const bool lifted_prev = r_lifted != 0;
    const bool lifted = (z_lift > p.lifting_bonus_threshold) || lifted_prev;
    const float lift_bonus_rew = p.lifting_bonus * ((lifted && !lifted_prev) ? 1.0f : 0.0f);
    const float not_lifted = lifted ? 0.0f : 1.0f;
    lifting_rew *= not_lifted;
    // _distance_delta_rewards (:759-786)
    float ft = 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float delta = ctip[i] - cur[i];
      ctip[i] = fminf(ctip[i], cur[i]);
      ft += clipf(delta, 0.0f, 10.0f) * coef[i];
    }
    const float hand_delta = furthest - cur[0];
    furthest = fmaxf(furthest, cur[0]);
    ft *= not_lifted;
    float hp = clipf(hand_delta, -10.0f, 0.0f);
    hp *= not_lifted;
    hp *= (float)kT1;
    // _keypoint_reward (:814-828)
    float kp_rew = clipf(ckp - kp_max, 0.0f, 100.0f);
    ckp = fminf(ckp, kp_max);
    kp_rew *= lifted ? 1.0f : 0.0f;
    // compute_kuka_reward (:854-930)
    const bool near_goal = kp_max <= keypoint_tol;
    const int near_steps = r_near + (near_goal ? 1 : 0);
    const bool is_success = near_steps >= p.success_steps;
    const float succ = r_succ + (is_success ? 1.0f : 0.0f);
    float t[HA_KUKA_N_EPISODE_TERMS];
    t[HA_KUKA_RAW_FINGERTIP_DELTA] = ft; t[HA_KUKA_RAW_HAND_DELTA] = hp; t[HA_KUKA_RAW_LIFTING] = lifting_rew; t[HA_KUKA_RAW_KEYPOINT] = kp_rew;
    ft *= p.distance_delta_rew_scale;
    hp *= p.hand_delta_rew_scale;
    lifting_rew *= p.lifting_rew_scale;
    kp_rew *= p.keypoint_rew_scale;
    const float kuka_pen = -1.0f * (sa * p.kuka_actions_penalty_scale);
    const float allegro_pen = -1.0f * (sh * p.allegro_actions_penalty_scale);
    const float bonus_rew = (near_goal ? 1.0f : 0.0f) * p.bonus_per_step;
    const float reward = ft + hp + lifting_rew + lift_bonus_rew + kp_rew + kuka_pen + allegro_pen + bonus_rew;
    t[HA_KUKA_FINGERTIP_DELTA] = ft; t[HA_KUKA_HAND_DELTA] = hp; t[HA_KUKA_LIFTING] = lifting_rew; t[HA_KUKA_LIFT_BONUS] = lift_bonus_rew;
    t[HA_KUKA_KEYPOINT] = kp_rew; t[HA_KUKA_BONUS] = bonus_rew; t[HA_KUKA_KUKA_PENALTY] = kuka_pen; t[HA_KUKA_ALLEGRO_PENALTY] = allegro_pen;
    // _compute_resets (:841-849) + reorientation _extra_reset_rules (allegro_kuka_reorientation.py:97-102)
    long long resets = (obj_z < p.fall_z) ? 1 : r_reset;
    long long prog = progress;
    if (p.max_consecutive_successes > 0) {
      prog = is_success ? 0 : prog;
      resets = (succ >= (float)p.max_consecutive_successes) ? 1 : resets;
    }
    const bool timed_out = prog >= (long long)(p.max_episode_length - 1);
    resets = timed_out ? 1 : resets;
    if (p.subtask == HA_KUKA_REORIENTATION) {
      const float far = fmaxf(fmaxf(cur[0], cur[1]), fmaxf(cur[2], cur[3]));
      resets = (far > p.far_tip_dist) ? 1 : resets;
    }
    const float true_obj = succ * tobj_scale + tobj_offset;
    const bool timeout_flag = timed_out && (resets != 0);          // vec_task.py:424
    o[L.reward] = reward * 0.01f;                                  // (:1437-1439)

    {
      KukaPostEnvStats st;
      st.valid = valid;
      st.reward = reward; st.prev_ep = r_prev_ep; st.tobj = true_obj;
      st.resets = resets != 0; st.success = is_success; st.timeout = timeout_flag; st.lifted = lifted;
      kuka_post_stats(b, st, lane);
    }
    // per-env outputs (lane = env: coalesced)
    if (valid) {
      b.rew_buf[env] = reward;
      b.reset_buf[env] = resets;
      b.progress_buf[env] = prog;
      b.reset_goal_buf[env] = is_success ? 1 : 0;
      if (b.randomize_buf) b.randomize_buf[env] = r_randomize + 1;   // (:1430)
      b.timeout_buf[env] = timeout_flag ? 1 : 0;
      b.successes[env] = succ;
      b.true_objective[env] = true_obj;
      b.near_goal_steps[env] = near_steps;
      b.lifted_object[env] = lifted ? 1 : 0;
      b.closest_keypoint_max_dist[env] = ckp;
      b.furthest_hand_dist[env] = furthest;
      *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
      // rewards_episode[k] += term (:871-874, :923-926)
#pragma unroll
      for (int k = 0; k < HA_KUKA_N_EPISODE_TERMS; ++k) b.rewards_episode[(size_t)k * n + env] = r_acc[k] + t[k];
      // episode_cumulative[k] = term, in the reference's list order (:912-921)
      constexpr int cum_src[HA_KUKA_N_STEP_TERMS] = {HA_KUKA_FINGERTIP_DELTA, HA_KUKA_HAND_DELTA, HA_KUKA_LIFTING, HA_KUKA_LIFT_BONUS,
                                                     HA_KUKA_KEYPOINT, HA_KUKA_KUKA_PENALTY, HA_KUKA_ALLEGRO_PENALTY, HA_KUKA_BONUS};
#pragma unroll
      for (int k = 0; k < HA_KUKA_N_STEP_TERMS; ++k) b.episode_cumulative[(size_t)k * n + env] = t[cum_src[k]];
    }
  }
  __syncwarp();
  {
    // obs tile: clamp (:1174-1176 clamp_obs, vec_task.py:432 clip) and store as one contiguous run
    const float c = p.clamp_abs_obs;
    const bool do_clamp = c > 0.0f;
    float4* dst = reinterpret_cast<float4*>(b.obs_buf + (size_t)env0 * OBS);   // 16-byte aligned: env0 % 32 == 0
    const float4* src = reinterpret_cast<const float4*>(s.obs);
    const int total = nvalid * OBS, nvec = total >> 2;
    constexpr int kObsIters = (kTile * OBS / 4 + 31) / 32;
#pragma unroll 6
    for (int i = 0; i < kObsIters; ++i) {
      const int q = lane + 32 * i;
      if (q < nvec) {
        float4 v = src[q];
        if (do_clamp) {
          v.x = fminf(fmaxf(v.x, -c), c); v.y = fminf(fmaxf(v.y, -c), c);
          v.z = fminf(fmaxf(v.z, -c), c); v.w = fminf(fmaxf(v.w, -c), c);
        }
        dst[q] = v;
      }
    }
    if (!kFull) {                                     // ragged last tile: up to 3 trailing floats
      const int f = (nvec << 2) + lane;
      if (f < total) {
        float v = s.obs[f];
        if (do_clamp) v = fminf(fmaxf(v, -c), c);
        b.obs_buf[(size_t)env0 * OBS + f] = v;
      }
    }
  }
}

template <int K>
__global__ void __launch_bounds__(kPostThreads)
kuka_post_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const int mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  KukaPostSmem<K>& sm = *reinterpret_cast<KukaPostSmem<K>*>(smem_raw);
  const int lane = threadIdx.x;
  const int env0 = blockIdx.x * kTile;
  const int nvalid = min(kTile, p.n_envs - env0);
  if (nvalid == kTile) kuka_post_tile<K, true>(p, b, sm, mode, env0, nvalid, lane);
  else kuka_post_tile<K, false>(p, b, sm, mode, env0, nvalid, lane);
}

}  // namespace ha
