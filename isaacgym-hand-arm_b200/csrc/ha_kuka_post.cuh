// K_post for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1426-1443 (post_physics_step) = compute_observations
// (:991-1089) + compute_full_state (:1091-1172) + compute_kuka_reward (:854-930)
// + _compute_resets (:841-849) + clamp_obs (:1174-1176), plus the VecTask.step
// tail (vec_task.py:424-432: timeout_buf, obs clip).
//
// Mapping (memory-bound design, no CTA barrier on the data path):
//   one WARP owns a tile of 8 consecutive envs and runs it start to finish;
//   lane = part * 8 + e  (e = env in tile, part = fingertip / keypoint index),
//   i.e. 4 lanes per env so that 32768 envs give 4096 independent warps.
//   phase L  the warp gathers the tile: contiguous spans (DOF state, keypoint
//            table, scales) as wide loads, AoS rows (13-float rigid-body /
//            root rows, 52-byte stride, only 4-byte aligned) with 16 or 8
//            lanes per row so one request touches 2..4 rows; everything is
//            issued before anything is consumed;
//   phase C  lane (part, e): fingertip `part` and keypoint `part` of env e;
//            the 4 lanes of an env exchange distances with shuffles and all
//            run the (identical) reward / reset logic;
//   phase S  per-env outputs are split over the 4 parts (coalesced 32..64-byte
//            runs); the 8 x obs_dim tile leaves as one contiguous run of
//            128-bit stores.
// Episode statistics: warp shuffle -> CTA -> 64-bit fixed-point atomics, so
// sums are order-independent (bit-reproducible); the last CTA converts them.
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
constexpr int kT1 = 4;               // fingertips == lanes per env
constexpr int kTile = 8;             // envs per warp
constexpr int kWarps = kThreads / 32;
constexpr int kNPart = HA_KUKA_N_PARTIAL_STATS;

// fixed-point scales of the 64-bit statistic accumulators
__host__ __device__ constexpr float stat_scale(int i) {
  return (i == HA_KUKA_STAT_SUM_REWARD) ? 65536.0f : (i == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE) ? 1048576.0f : 1.0f;
}
__host__ __device__ constexpr bool stat_is_max(int i) {
  return i == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE || i == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE;
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
struct KukaPostWarpSmem {
  static constexpr int kObs = kuka_layout(kD1, kT1, K).dim;
  float obs[kTile * kObs];          // row-major tile, leaves as one contiguous run
  float palm[kTile * 13];
  float tip[kTile * kT1 * 7];       // [(e * 4 + t) * 7 + c]: pos3 quat4
  float obj[kTile * 13];
  float goal[kTile * 7];
  float kp[kTile * 3 * K];
  float lo[kD1 + 1], hi[kD1 + 1];   // per-DOF limits (no divergent constant-bank reads in the loops)
};

template <int K>
struct KukaPostSmem {
  KukaPostWarpSmem<K> w[kWarps];
  float part[kWarps][kNPart];
  int last_flag;
};

struct KukaPostEnvStats {   // per-env results that feed the statistics (idle lanes: valid == false)
  float reward = 0.f, resets = 0.f, success = 0.f, prev_ep = 0.f, tobj = 0.f, timeout = 0.f, lifted = 0.f;
  bool valid = false;
};

// One warp, one tile of up to 8 envs.  kFull: all 8 envs exist (no index clamps; every warp but the last one).
template <int K, bool kFull>
__device__ __forceinline__ void kuka_post_tile(const HaKukaParams& p, const HaKukaBuffers& b, KukaPostWarpSmem<K>& s, const int mode,
                                               const int env0, const int nvalid_rt, const int lane, KukaPostEnvStats& st) {
  using WS = KukaPostWarpSmem<K>;
  constexpr int OBS = WS::kObs;
  constexpr KukaLayout L = kuka_layout(kD1, kT1, K);
  const int n = p.n_envs;
  const int nvalid = kFull ? kTile : nvalid_rt;
  const int part = lane >> 3, el = lane & 7;        // compute mapping: 4 lanes per env
  const bool valid = kFull ? true : (el < nvalid);
  const unsigned FULL = 0xffffffffu;
  const int last = nvalid - 1;
  auto cl = [&](int e) { return kFull ? e : min(e, last); };   // idle lanes of a ragged tile re-read a valid env
  {
    const int env = env0 + cl(el);                  // stores are predicated on `valid`
    // ------------------------------ phase L ---------------------------------
    if (lane < kD1) { s.lo[lane] = p.dof_lower[lane]; s.hi[lane] = p.dof_upper[lane]; }

    // AoS rows.  16 lanes per 13-float row (2 rows / request), 8 lanes per 7-float row (4 rows / request).
    const int h16 = lane >> 4, c16 = lane & 15;
    const int r8 = lane >> 3, c8 = lane & 7;
    const size_t rb_env_stride = (size_t)p.n_bodies * 13;
    const float* rb0 = b.rigid_body_state + (size_t)env0 * rb_env_stride;
    float palm_v[4], obj_v[4], tip_v[8], goal_v[2];
    {
      const float* src = rb0 + (size_t)p.palm_body[0] * 13 + min(c16, 12);
#pragma unroll
      for (int i = 0; i < 4; ++i) palm_v[i] = ldg(src + (size_t)cl(2 * i + h16) * rb_env_stride);
    }
    {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int e = env0 + cl(2 * i + h16);
        const int row = p.root_regular ? e * p.actors_per_env + p.object_slot : __ldg(b.object_actor + e);
        obj_v[i] = b.root_state[(size_t)row * 13 + min(c16, 12)];
      }
    }
    {
      const float* src = rb0 + (size_t)p.tip_body[r8] * 13 + min(c8, 6);
#pragma unroll
      for (int i = 0; i < 8; ++i) tip_v[i] = ldg(src + (size_t)cl(i) * rb_env_stride);
    }
    {
#pragma unroll
      for (int i = 0; i < 2; ++i) goal_v[i] = b.goal_states[(size_t)(env0 + cl(4 * i + r8)) * 13 + min(c8, 6)];
    }
    // contiguous spans: DOF state [nvalid, 23, 2], keypoint table [nvalid, K, 3], scales [nvalid, 3]
    const float* dof_src = b.dof_state + (size_t)env0 * (2 * kD1);
    const int dof_total = nvalid * 2 * kD1, dof_nvec = dof_total >> 2;
    float4 dof_v[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int q = lane + 32 * i;
      dof_v[i] = (q < dof_nvec) ? __ldg(reinterpret_cast<const float4*>(dof_src) + q) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    float dof_tail = 0.f;                             // odd nvalid leaves 2 floats after the last float4
    if ((dof_nvec << 2) + lane < dof_total) dof_tail = ldg(dof_src + (dof_nvec << 2) + lane);
    constexpr int kKpIters = (kTile * 3 * K + 31) / 32;
    float kp_v[kKpIters];
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = lane + 32 * i;
      kp_v[i] = (idx < nvalid * 3 * K) ? ldg(b.object_keypoint_offsets + (size_t)env0 * 3 * K + idx) : 0.f;
    }
    const float scale_v = (lane < nvalid * 3) ? ldg(b.object_scales + (size_t)env0 * 3 + lane) : 0.f;

    // per-env scalars (all 4 lanes of an env read the same words: one sector per request)
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
    float r_acc[3] = {0.f, 0.f, 0.f};                  // episode accumulators: part p owns terms p, p+4, p+8
    long long r_randomize = 0;
    if (mode == kPostFull) {
#pragma unroll
      for (int j = 0; j < 3; ++j) r_acc[j] = b.rewards_episode[(size_t)(part + 4 * j) * n + env];
      if (b.randomize_buf && part == 1) r_randomize = b.randomize_buf[env];
    }
    const float keypoint_tol = b.dynamic->keypoint_success_tolerance;
    const float tobj_scale = b.dynamic->true_objective_scale, tobj_offset = b.dynamic->true_objective_offset;

    // ---- all loads issued; scatter into the warp's smem ----
    __syncwarp();                                     // lo / hi visible
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = 2 * i + h16;
      if (c16 < 13) {
        s.palm[e * 13 + c16] = palm_v[i];
        s.obj[e * 13 + c16] = obj_v[i];
        if (c16 >= 3) {                               // palm rot + vel (:1112-1117), object rot + vel (:1120)
          s.obs[e * OBS + L.palm_rot_vel + c16 - 3] = palm_v[i];
          s.obs[e * OBS + L.obj_rot_vel + c16 - 3] = obj_v[i];
        }
      }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i)
      if (c8 < 7) s.tip[(i * kT1 + r8) * 7 + c8] = tip_v[i];
#pragma unroll
    for (int i = 0; i < 2; ++i)
      if (c8 < 7) s.goal[(4 * i + r8) * 7 + c8] = goal_v[i];
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = lane + 32 * i;
      if (idx < kTile * 3 * K) s.kp[idx] = kp_v[i];
    }
    if (lane < kTile * 3) { const int e = lane / 3; s.obs[e * OBS + L.scales + (lane - 3 * e)] = scale_v; }   // (:1143)
    // DOF part of the obs row: unscale(dof_pos) (:1095-1100), raw dof_vel (:1104)
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      const int q = lane + 32 * i;
      if (q < dof_nvec) {
        const int g0 = 2 * q;                         // (env, dof) pair index; a float4 holds pairs g0 and g0 + 1
        const int e0 = g0 / kD1, d0 = g0 - e0 * kD1;
        const int e1 = (d0 == kD1 - 1) ? e0 + 1 : e0, d1 = (d0 == kD1 - 1) ? 0 : d0 + 1;
        s.obs[e0 * OBS + L.dof_pos + d0] = unscale_from(dof_v[i].x, s.lo[d0], s.hi[d0]);
        s.obs[e0 * OBS + L.dof_vel + d0] = dof_v[i].y;
        s.obs[e1 * OBS + L.dof_pos + d1] = unscale_from(dof_v[i].z, s.lo[d1], s.hi[d1]);
        s.obs[e1 * OBS + L.dof_vel + d1] = dof_v[i].w;
      }
    }
    if ((dof_nvec << 2) + lane < dof_total) {         // the last (pos, vel) pair of an odd tile
      const int f = (dof_nvec << 2) + lane;
      const int e = f / (2 * kD1), r = f - e * (2 * kD1), d = r >> 1;
      s.obs[e * OBS + ((r & 1) ? L.dof_vel + d : L.dof_pos + d)] = (r & 1) ? dof_tail : unscale_from(dof_tail, s.lo[d], s.hi[d]);
    }
    __syncwarp();

    // ------------------------------ phase C ---------------------------------
    float* o = s.obs + el * OBS;
    const float* pr = s.palm + el * 13;
    const float* ob = s.obj + el * 13;
    // palm_center_pos = palm_pos + quat_rotate(palm_rot, palm_offset)   (:1013-1017)
    const f3 pc = f3{pr[0], pr[1], pr[2]} + quat_rotate(f4{pr[3], pr[4], pr[5], pr[6]}, f3{p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]});
    const f3 obj_pos = {ob[0], ob[1], ob[2]};
    const float obj_z = ob[2];
    float my_tip_dist;
    {   // fingertip `part` (:1033-1052)
      const float* tr = s.tip + (el * kT1 + part) * 7;
      const f3 to = f3{tr[0], tr[1], tr[2]} +
                    quat_rotate(f4{tr[3], tr[4], tr[5], tr[6]}, f3{p.tip_offset[part][0], p.tip_offset[part][1], p.tip_offset[part][2]});
      const f3 rel_obj = to - obj_pos, rel_palm = to - pc;
      my_tip_dist = norm3(rel_obj);
      o[L.tip_rel_palm + 3 * part + 0] = rel_palm.x;
      o[L.tip_rel_palm + 3 * part + 1] = rel_palm.y;
      o[L.tip_rel_palm + 3 * part + 2] = rel_palm.z;
    }
    float my_kp_dist = 0.0f;                          // distances are >= 0, so 0 is neutral for the max
    if (part < K) {                                   // keypoint `part` (:1057-1073)
      const float* kp = s.kp + el * 3 * K + 3 * part;
      const float* gl = s.goal + el * 7;
      const f3 off = {kp[0], kp[1], kp[2]};
      const f3 okp = obj_pos + quat_rotate(f4{ob[3], ob[4], ob[5], ob[6]}, off);
      const f3 gkp = f3{gl[0], gl[1], gl[2]} + quat_rotate(f4{gl[3], gl[4], gl[5], gl[6]}, off);
      const f3 rel_goal = okp - gkp, rel_palm = okp - pc;
      my_kp_dist = norm3(rel_goal);
      o[L.kp_rel_palm + 3 * part + 0] = rel_palm.x;
      o[L.kp_rel_palm + 3 * part + 1] = rel_palm.y;
      o[L.kp_rel_palm + 3 * part + 2] = rel_palm.z;
      o[L.kp_rel_goal + 3 * part + 0] = rel_goal.x;
      o[L.kp_rel_goal + 3 * part + 1] = rel_goal.y;
      o[L.kp_rel_goal + 3 * part + 2] = rel_goal.z;
    }
    // _action_penalties sums (:830-839): each part adds its share, then a 4-lane butterfly
    float sa = 0.f, sh = 0.f;
    {
      const float* v = o + L.dof_vel;
      sa = fabsf(v[part]);
      if (part < 3) sa += fabsf(v[part + 4]);
#pragma unroll
      for (int j = 0; j < 4; ++j) sh += fabsf(v[7 + part + 4 * j]);
    }
    sa += __shfl_xor_sync(FULL, sa, 8);  sh += __shfl_xor_sync(FULL, sh, 8);
    sa += __shfl_xor_sync(FULL, sa, 16); sh += __shfl_xor_sync(FULL, sh, 16);
    float cur[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) cur[j] = __shfl_sync(FULL, my_tip_dist, j * 8 + el);
    float kp_max = fmaxf(my_kp_dist, __shfl_xor_sync(FULL, my_kp_dist, 8));
    kp_max = fmaxf(kp_max, __shfl_xor_sync(FULL, kp_max, 16));

    // ------------------------ reward / reset logic ---------------------------
    // (identical in the 4 lanes of an env; each part stores its share of the results)
    float ctip[4] = {r_ctip.x, r_ctip.y, r_ctip.z, r_ctip.w};
    const float coef[4] = {r_coef.x, r_coef.y, r_coef.z, r_coef.w};
    // lazy initialisation of the per-episode closest / furthest distances (:1044-1049, :1077-1079)
#pragma unroll
    for (int i = 0; i < 4; ++i) ctip[i] = (ctip[i] < 0.0f) ? cur[i] : ctip[i];
    float furthest = (r_furthest < 0.0f) ? cur[0] : r_furthest;
    float ckp = (r_ckp < 0.0f) ? kp_max : r_ckp;
    const long long progress = r_progress + (mode == kPostFull ? 1 : 0);   // progress_buf += 1 (:1429)
    // observation slots that see pre-update state (ordering: SURVEY 8a)
    if (part == 0) {
      o[L.palm_pos + 0] = pc.x; o[L.palm_pos + 1] = pc.y; o[L.palm_pos + 2] = pc.z;
      o[L.closest_kp] = ckp;
      o[L.lifted] = r_lifted ? 1.0f : 0.0f;
    }
    o[L.closest_tip + part] = (part == 0) ? ctip[0] : (part == 1) ? ctip[1] : (part == 2) ? ctip[2] : ctip[3];
    {   // log(progress / 10 + 1) (:1162) by part 0, log(successes + 1) (:1164) by part 1: one logf per lane
      const float lg = logf((part == 0) ? ((float)progress / 10.0f + 1.0f) : (r_succ + 1.0f));
      if (part < 2) o[(part == 0) ? L.progress : L.successes] = lg;
    }

    if (mode == kPostObserveOnly) {
      if (part == 1) o[L.reward] = r_old_rew_obs;
      if (valid && part == 3) {
        b.closest_keypoint_max_dist[env] = ckp;
        b.furthest_hand_dist[env] = furthest;
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
      }
    } else {
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
      if (part == 1) o[L.reward] = reward * 0.01f;                   // (:1437-1439)

      // ------------------------------ phase S ---------------------------------
      // per-env outputs, split over the 4 parts of an env
      if (valid) {
        if (part == 0) {
          b.rew_buf[env] = reward;
          b.reset_buf[env] = resets;
          b.progress_buf[env] = prog;
        } else if (part == 1) {
          b.reset_goal_buf[env] = is_success ? 1 : 0;
          if (b.randomize_buf) b.randomize_buf[env] = r_randomize + 1;   // (:1430)
          b.timeout_buf[env] = timeout_flag ? 1 : 0;
        } else if (part == 2) {
          b.successes[env] = succ;
          b.true_objective[env] = true_obj;
          b.near_goal_steps[env] = near_steps;
          b.lifted_object[env] = lifted ? 1 : 0;
        } else {
          b.closest_keypoint_max_dist[env] = ckp;
          b.furthest_hand_dist[env] = furthest;
          *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
        }
        // rewards_episode[k] += term (:871-874, :923-926): part p owns k = p, p + 4, p + 8
#pragma unroll
        for (int j = 0; j < 3; ++j) {
          const float tk = (part == 0) ? t[4 * j] : (part == 1) ? t[4 * j + 1] : (part == 2) ? t[4 * j + 2] : t[4 * j + 3];
          b.rewards_episode[(size_t)(part + 4 * j) * n + env] = r_acc[j] + tk;
        }
        // episode_cumulative[k] = term, the reference's list order (:912-921): part p owns k = p, p + 4
        {
          const float c0 = (part == 0) ? t[HA_KUKA_FINGERTIP_DELTA] : (part == 1) ? t[HA_KUKA_HAND_DELTA]
                         : (part == 2) ? t[HA_KUKA_LIFTING] : t[HA_KUKA_LIFT_BONUS];
          const float c1 = (part == 0) ? t[HA_KUKA_KEYPOINT] : (part == 1) ? t[HA_KUKA_KUKA_PENALTY]
                         : (part == 2) ? t[HA_KUKA_ALLEGRO_PENALTY] : t[HA_KUKA_BONUS];
          b.episode_cumulative[(size_t)part * n + env] = c0;
          b.episode_cumulative[(size_t)(part + 4) * n + env] = c1;
        }
      }
      if (valid) {
        st.valid = true;
        st.reward = reward; st.resets = (float)resets; st.success = is_success ? 1.f : 0.f; st.prev_ep = r_prev_ep;
        st.tobj = true_obj; st.timeout = timeout_flag ? 1.f : 0.f; st.lifted = lifted ? 1.f : 0.f;
      }
    }
    __syncwarp();
    {
      // obs tile: clamp (:1174-1176 clamp_obs, vec_task.py:432 clip) and store as one contiguous run
      const float c = p.clamp_abs_obs;
      const bool do_clamp = c > 0.0f;
      float* dst = b.obs_buf + (size_t)env0 * OBS;
      const int total = nvalid * OBS;
      const int nvec = total >> 2;                    // tile base is 16-byte aligned (env0 % 8 == 0, obs_buf aligned)
      constexpr int kObsIters = (kTile * OBS / 4 + 31) / 32;
#pragma unroll
      for (int i = 0; i < kObsIters; ++i) {
        const int q = lane + 32 * i;
        if (q < nvec) {
          float4 v = *reinterpret_cast<const float4*>(s.obs + 4 * q);
          if (do_clamp) {
            v.x = fminf(fmaxf(v.x, -c), c); v.y = fminf(fmaxf(v.y, -c), c);
            v.z = fminf(fmaxf(v.z, -c), c); v.w = fminf(fmaxf(v.w, -c), c);
          }
          *reinterpret_cast<float4*>(dst + 4 * q) = v;
        }
      }
      const int f = (nvec << 2) + lane;               // ragged last tile: up to 3 trailing floats
      if (f < total) {
        float v = s.obs[f];
        if (do_clamp) v = fminf(fmaxf(v, -c), c);
        dst[f] = v;
      }
    }
  }
}

template <int K>
__global__ void __launch_bounds__(kThreads, 7)
kuka_post_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const int mode) {
  using Smem = KukaPostSmem<K>;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& sm = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int env0 = (blockIdx.x * kWarps + warp) * kTile;
  const int nvalid = min(kTile, p.n_envs - env0);          // <= 0: this warp has no tile
  const int part = lane >> 3;
  const unsigned FULL = 0xffffffffu;
  KukaPostEnvStats st;
  if (nvalid == kTile) kuka_post_tile<K, true>(p, b, sm.w[warp], mode, env0, nvalid, lane, st);
  else if (nvalid > 0) kuka_post_tile<K, false>(p, b, sm.w[warp], mode, env0, nvalid, lane, st);
  if (mode != kPostFull) return;

  // --------------------------- episode statistics ---------------------------
  // warp: one lane per env contributes (part 0); 0/1 flags by ballot, the rest by an 8-lane butterfly
  {
    const bool c = st.valid && part == 0;
    const unsigned m_n = __ballot_sync(FULL, c), m_rs = __ballot_sync(FULL, c && st.resets != 0.f);
    const unsigned m_su = __ballot_sync(FULL, c && st.success != 0.f), m_to = __ballot_sync(FULL, c && st.timeout != 0.f);
    const unsigned m_li = __ballot_sync(FULL, c && st.lifted != 0.f);
    float v_rew = c ? st.reward : 0.f, v_pe = c ? st.prev_ep : 0.f, v_to = c ? st.tobj : 0.f;
    float v_mx = c ? st.tobj : -3.0e38f, v_mn = c ? -st.tobj : -3.0e38f;
#pragma unroll
    for (int ofs = 4; ofs > 0; ofs >>= 1) {
      v_rew += __shfl_xor_sync(FULL, v_rew, ofs);
      v_pe += __shfl_xor_sync(FULL, v_pe, ofs);
      v_to += __shfl_xor_sync(FULL, v_to, ofs);
      v_mx = fmaxf(v_mx, __shfl_xor_sync(FULL, v_mx, ofs));
      v_mn = fmaxf(v_mn, __shfl_xor_sync(FULL, v_mn, ofs));
    }
    if (lane == 0) {
      float* o = sm.part[warp];
      o[HA_KUKA_STAT_N] = (float)__popc(m_n);
      o[HA_KUKA_STAT_SUM_REWARD] = v_rew;
      o[HA_KUKA_STAT_SUM_RESETS] = (float)__popc(m_rs);
      o[HA_KUKA_STAT_SUM_SUCCESS_EVENTS] = (float)__popc(m_su);
      o[HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES] = v_pe;
      o[HA_KUKA_STAT_SUM_TRUE_OBJECTIVE] = v_to;
      o[HA_KUKA_STAT_MAX_TRUE_OBJECTIVE] = v_mx;
      o[HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE] = v_mn;
      o[HA_KUKA_STAT_SUM_TIMEOUTS] = (float)__popc(m_to);
      o[HA_KUKA_STAT_SUM_LIFTED] = (float)__popc(m_li);
    }
  }
  __syncthreads();
  if (tid < kNPart) {
    float acc = sm.part[0][tid];
#pragma unroll
    for (int w = 1; w < kWarps; ++w) acc = stat_is_max(tid) ? fmaxf(acc, sm.part[w][tid]) : acc + sm.part[w][tid];
    if (stat_is_max(tid)) {
      atomicMax(reinterpret_cast<long long*>(b.stat_accum + tid), float_key(acc));
    } else {
      const float sc = (tid == HA_KUKA_STAT_SUM_REWARD) ? 65536.0f : (tid == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE) ? 1048576.0f : 1.0f;
      atomicAdd(reinterpret_cast<unsigned long long*>(b.stat_accum + tid), (unsigned long long)__float2ll_rn(acc * sc));
    }
  }
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    const unsigned tk = atomicAdd(b.tickets + 0, 1u);
    sm.last_flag = (tk == gridDim.x - 1);
  }
  __syncthreads();
  if (!sm.last_flag) return;
  // ---------------------- last CTA: final statistics ----------------------
  if (tid == 0) __threadfence();
  if (tid < kNPart) {
    const bool is_max = stat_is_max(tid);
    const long long raw = (long long)atomicExch(reinterpret_cast<unsigned long long*>(b.stat_accum + tid),
                                                (unsigned long long)(is_max ? kStatMaxInit : 0LL));
    const float sc = (tid == HA_KUKA_STAT_SUM_REWARD) ? 65536.0f : (tid == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE) ? 1048576.0f : 1.0f;
    const float val = is_max ? key_float(raw) : (float)((double)raw / (double)sc);
    b.stats[tid] = val;
    sm.part[0][tid] = val;
  }
  __syncthreads();
  if (tid == 0) {
    // means / min for extras (allegro_kuka_base.py:903-910), so the host needs no extra launches
    const float inv_n = 1.0f / sm.part[0][HA_KUKA_STAT_N];
    b.stats[HA_KUKA_STAT_MEAN_REWARD] = sm.part[0][HA_KUKA_STAT_SUM_REWARD] * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES] = sm.part[0][HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES] * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE] = sm.part[0][HA_KUKA_STAT_SUM_TRUE_OBJECTIVE] * inv_n;
    b.stats[HA_KUKA_STAT_MIN_TRUE_OBJECTIVE] = -sm.part[0][HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE];
    b.tickets[0] = 0;                    // re-arm for the next launch (graph replay safe)
    // the reset/index counters of the pre-step are consumed by now: clear for the next pre-step
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    // fast-mode RNG: one counter tick per step
    b.dynamic->rng_step = b.dynamic->rng_step + 1;
  }
}

}  // namespace ha
