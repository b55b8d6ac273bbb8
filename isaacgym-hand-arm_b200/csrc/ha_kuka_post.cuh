// K_post for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1426-1443 (post_physics_step) = compute_observations
// (:991-1089) + compute_full_state (:1091-1172) + compute_kuka_reward (:854-930)
// + _compute_resets (:841-849) + clamp_obs (:1174-1176), plus the VecTask.step
// tail (vec_task.py:424-432: timeout_buf, obs clip).
//
// Mapping: one CTA = a tile of 32 consecutive envs, 128 threads.
//   phase L  all threads gather the tile's inputs with coalesced loads
//            (contiguous spans for DOF / keypoint tables, 28..52-byte row
//            granules for the AoS rigid-body / root rows) into padded smem;
//   phase C  warp w, lane e: fingertip w and keypoint w of env e (+ one of the
//            per-env odd jobs); results go to the smem obs tile;
//   phase R  warp 0, lane e: reward / success / reset logic of env e;
//   phase S  all threads: episode accumulators (SoA, coalesced) and the obs
//            tile as one contiguous run of 128-bit stores.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"

namespace ha {

struct KukaLayout {            // obs slot offsets, allegro_kuka_base.py:1091-1172
  int dof_pos, dof_vel, palm_pos, palm_rot_vel, obj_rot_vel, tip_rel_palm, kp_rel_palm, kp_rel_goal,
      scales, closest_kp, closest_tip, lifted, progress, successes, reward, dim;
};
__host__ __device__ inline KukaLayout kuka_layout(int n_dofs, int n_tips, int K) {
  KukaLayout l;
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
constexpr int kRbW = 13 + 7 * kT1;   // palm row + (pos,quat) of 4 tips = 41 (odd: conflict-free)
constexpr int kRedW = 13;            // per-env reduction scratch row (odd)
constexpr int kNPart = HA_KUKA_N_PARTIAL_STATS;

template <int K>
struct KukaPostSmem {
  static constexpr int kObs = 2 * kD1 + 3 + 10 + 10 + 3 * kT1 + 6 * K + 3 + 1 + kT1 + 1 + 1 + 1 + 1;
  float obs[kEnvsPerCta * kObs];     // flat row-major tile, stored contiguously
  float rb[kEnvsPerCta * kRbW];
  float obj[kEnvsPerCta * 13];
  float goal[kEnvsPerCta * 7];
  float kp[kEnvsPerCta * (3 * K | 1)];
  float red[kEnvsPerCta * kRedW];    // [0:4] tip dist, [4:8] kp dist, [8] sum|v arm|, [9] sum|v hand|
  float terms[HA_KUKA_N_EPISODE_TERMS * kEnvsPerCta];
  float part[4 * 16];
  int last_flag;
};

template <int K>
__global__ void __launch_bounds__(kThreads, 8)
kuka_post_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const int mode) {
  using Smem = KukaPostSmem<K>;
  constexpr int OBS = Smem::kObs;
  constexpr int KPW = (3 * K) | 1;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int env0 = blockIdx.x * kEnvsPerCta;
  const int n = p.n_envs;
  const int nvalid = min(kEnvsPerCta, n - env0);
  const KukaLayout L = kuka_layout(kD1, kT1, K);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;

  // ------------------------------ phase L ---------------------------------
  // per-env scalars of the reward phase go straight to registers of warp 0;
  // issue them first so they are in flight during the tile gather.
  long long r_progress = 0, r_reset = 0;
  float r_succ = 0.f, r_ckp = 0.f, r_furthest = 0.f, r_initz = 0.f, r_prev_ep = 0.f, r_old_rew_obs = 0.f;
  int r_near = 0;
  unsigned char r_lifted = 0;
  float4 r_ctip = make_float4(0.f, 0.f, 0.f, 0.f), r_coef = make_float4(1.f, 1.f, 1.f, 1.f);
  if (warp == 0 && valid) {
    r_progress = b.progress_buf[env];
    r_reset = b.reset_buf[env];
    r_succ = b.successes[env];
    r_ckp = b.closest_keypoint_max_dist[env];
    r_furthest = b.furthest_hand_dist[env];
    r_near = b.near_goal_steps[env];
    r_lifted = b.lifted_object[env];
    r_initz = ldg(b.object_init_state + (size_t)env * 13 + 2);
    r_prev_ep = b.prev_episode_successes[env];
    r_ctip = *reinterpret_cast<const float4*>(b.closest_fingertip_dist + (size_t)env * 4);
    r_coef = __ldg(reinterpret_cast<const float4*>(b.finger_rew_coeffs + (size_t)env * 4));
    if (mode == kPostObserveOnly) r_old_rew_obs = b.obs_buf[(size_t)env * OBS + L.reward];
  }
  // episode accumulators: warp w owns terms w, w+4, w+8
  float r_acc[3] = {0.f, 0.f, 0.f};
  if (mode == kPostFull && valid) {
#pragma unroll
    for (int j = 0; j < 3; ++j) r_acc[j] = b.rewards_episode[(size_t)(warp + 4 * j) * n + env];
  }

  // DOF tile: [nvalid, 23, 2] contiguous, base 16-byte aligned when env0 % 2 == 0
  {
    const float* src = b.dof_state + (size_t)env0 * (2 * kD1);
    const int total = nvalid * 2 * kD1;
    constexpr int kVecIters = (kEnvsPerCta * 2 * kD1 / 4 + kThreads - 1) / kThreads;  // 3
    float4 v[kVecIters];
    const int nvec = total >> 2;
#pragma unroll
    for (int i = 0; i < kVecIters; ++i) {
      const int q = tid + i * kThreads;
      if (q < nvec) v[i] = __ldg(reinterpret_cast<const float4*>(src) + q);
    }
    // rigid-body rows: palm (13 floats) + 4 tips (7 floats each) per env
    constexpr int kRbIters = (kEnvsPerCta * kRbW + kThreads - 1) / kThreads;  // 11
    float rbv[kRbIters];
    const float* rb = b.rigid_body_state + (size_t)env0 * p.n_bodies * 13;
#pragma unroll
    for (int i = 0; i < kRbIters; ++i) {
      const int idx = tid + i * kThreads;
      const int e = idx / kRbW, j = idx - e * kRbW;
      int body, c;
      if (j < 13) { body = p.palm_body[0]; c = j; }
      else { const int t = (j - 13) / 7; c = (j - 13) - 7 * t; body = p.tip_body[t]; }
      rbv[i] = (e < nvalid) ? ldg(rb + ((size_t)e * p.n_bodies + body) * 13 + c) : 0.f;
    }
    // object root row (13) and goal pose (7)
    constexpr int kObjIters = (kEnvsPerCta * 13 + kThreads - 1) / kThreads;  // 4
    float ov[kObjIters];
#pragma unroll
    for (int i = 0; i < kObjIters; ++i) {
      const int idx = tid + i * kThreads;
      const int e = idx / 13, c = idx - e * 13;
      ov[i] = 0.f;
      if (e < nvalid) {
        const int row = __ldg(b.object_actor + env0 + e);
        ov[i] = b.root_state[(size_t)row * 13 + c];
      }
    }
    constexpr int kGoalIters = (kEnvsPerCta * 7 + kThreads - 1) / kThreads;  // 2
    float gv[kGoalIters];
#pragma unroll
    for (int i = 0; i < kGoalIters; ++i) {
      const int idx = tid + i * kThreads;
      const int e = idx / 7, c = idx - e * 7;
      gv[i] = (e < nvalid) ? b.goal_states[(size_t)(env0 + e) * 13 + c] : 0.f;
    }
    constexpr int kKpIters = (kEnvsPerCta * 3 * K + kThreads - 1) / kThreads;
    float kv[kKpIters];
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = tid + i * kThreads;
      kv[i] = (idx < nvalid * 3 * K) ? ldg(b.object_keypoint_offsets + (size_t)env0 * 3 * K + idx) : 0.f;
    }

    // ---- all loads issued; now scatter into smem ----
#pragma unroll
    for (int i = 0; i < kVecIters; ++i) {
      const int q = tid + i * kThreads;
      if (q < nvec) {
        const float vv[4] = {v[i].x, v[i].y, v[i].z, v[i].w};
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int f = 4 * q + c;
          const int e = f / (2 * kD1), r = f - e * (2 * kD1);
          const int d = r >> 1;
          // dof_pos -> unscale (allegro_kuka_base.py:1095-1100); dof_vel raw (:1104)
          s.obs[e * OBS + ((r & 1) ? L.dof_vel + d : L.dof_pos + d)] =
              (r & 1) ? vv[c] : unscale_from(vv[c], p.dof_lower[d], p.dof_upper[d]);
        }
      }
    }
    for (int f = (nvec << 2) + tid; f < total; f += kThreads) {  // tail of a ragged last tile
      const int e = f / (2 * kD1), r = f - e * (2 * kD1), d = r >> 1;
      const float x = src[f];
      s.obs[e * OBS + ((r & 1) ? L.dof_vel + d : L.dof_pos + d)] = (r & 1) ? x : unscale_from(x, p.dof_lower[d], p.dof_upper[d]);
    }
#pragma unroll
    for (int i = 0; i < kRbIters; ++i) {
      const int idx = tid + i * kThreads;
      if (idx < kEnvsPerCta * kRbW) s.rb[idx] = rbv[i];
    }
#pragma unroll
    for (int i = 0; i < kObjIters; ++i) {
      const int idx = tid + i * kThreads;
      if (idx < kEnvsPerCta * 13) s.obj[idx] = ov[i];
    }
#pragma unroll
    for (int i = 0; i < kGoalIters; ++i) {
      const int idx = tid + i * kThreads;
      if (idx < kEnvsPerCta * 7) s.goal[idx] = gv[i];
    }
#pragma unroll
    for (int i = 0; i < kKpIters; ++i) {
      const int idx = tid + i * kThreads;
      if (idx < kEnvsPerCta * 3 * K) { const int e = idx / (3 * K); s.kp[e * KPW + (idx - e * 3 * K)] = kv[i]; }
    }
  }
  __syncthreads();

  // ------------------------------ phase C ---------------------------------
  {
    const float* rb = s.rb + lane * kRbW;
    const float* ob = s.obj + lane * 13;
    float* o = s.obs + lane * OBS;
    const f3 palm_pos = {rb[0], rb[1], rb[2]};
    const f4 palm_rot = {rb[3], rb[4], rb[5], rb[6]};
    // palm_center_pos = palm_pos + quat_rotate(palm_rot, palm_offset)   (:1013-1017)
    const f3 pc = palm_pos + quat_rotate(palm_rot, f3{p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]});
    const f3 obj_pos = {ob[0], ob[1], ob[2]};
    // fingertip `warp` (:1033-1052)
    {
      const int t = warp;
      const float* tr = rb + 13 + 7 * t;
      const f3 tp = {tr[0], tr[1], tr[2]};
      const f4 tq = {tr[3], tr[4], tr[5], tr[6]};
      const f3 to = tp + quat_rotate(tq, f3{p.tip_offset[t][0], p.tip_offset[t][1], p.tip_offset[t][2]});
      const f3 rel_obj = to - obj_pos;
      const f3 rel_palm = to - pc;
      s.red[lane * kRedW + t] = norm3(rel_obj);
      o[L.tip_rel_palm + 3 * t + 0] = rel_palm.x;
      o[L.tip_rel_palm + 3 * t + 1] = rel_palm.y;
      o[L.tip_rel_palm + 3 * t + 2] = rel_palm.z;
    }
    // keypoint `warp` (:1057-1073)
    if (warp < K) {
      const int k = warp;
      const float* kp = s.kp + lane * KPW + 3 * k;
      const float* gl = s.goal + lane * 7;
      const f3 off = {kp[0], kp[1], kp[2]};
      const f4 obj_rot = {ob[3], ob[4], ob[5], ob[6]};
      const f3 okp = obj_pos + quat_rotate(obj_rot, off);
      const f3 gkp = f3{gl[0], gl[1], gl[2]} + quat_rotate(f4{gl[3], gl[4], gl[5], gl[6]}, off);
      const f3 rel_goal = okp - gkp;
      const f3 rel_palm = okp - pc;
      s.red[lane * kRedW + 4 + k] = norm3(rel_goal);
      o[L.kp_rel_palm + 3 * k + 0] = rel_palm.x;
      o[L.kp_rel_palm + 3 * k + 1] = rel_palm.y;
      o[L.kp_rel_palm + 3 * k + 2] = rel_palm.z;
      o[L.kp_rel_goal + 3 * k + 0] = rel_goal.x;
      o[L.kp_rel_goal + 3 * k + 1] = rel_goal.y;
      o[L.kp_rel_goal + 3 * k + 2] = rel_goal.z;
    }
    if (warp == 0) {          // palm pos, palm rot + vel (:1112-1117)
      o[L.palm_pos + 0] = pc.x; o[L.palm_pos + 1] = pc.y; o[L.palm_pos + 2] = pc.z;
#pragma unroll
      for (int c = 0; c < 10; ++c) o[L.palm_rot_vel + c] = rb[3 + c];
    } else if (warp == 1) {   // object rot + vel (:1120), object scales (:1143)
#pragma unroll
      for (int c = 0; c < 10; ++c) o[L.obj_rot_vel + c] = ob[3 + c];
#pragma unroll
      for (int c = 0; c < 3; ++c) o[L.scales + c] = valid ? ldg(b.object_scales + (size_t)env * 3 + c) : 0.f;
    } else if (warp == 2) {   // _action_penalties sums (:830-839)
      float sa = 0.f, sh = 0.f;
#pragma unroll
      for (int d = 0; d < 7; ++d) sa += fabsf(o[L.dof_vel + d]);
#pragma unroll
      for (int d = 7; d < kD1; ++d) sh += fabsf(o[L.dof_vel + d]);
      s.red[lane * kRedW + 8] = sa;
      s.red[lane * kRedW + 9] = sh;
    }
  }
  __syncthreads();

  // ------------------------------ phase R ---------------------------------
  if (warp == 0) {
    float* o = s.obs + lane * OBS;
    const float* rd = s.red + lane * kRedW;
    const float cur[4] = {rd[0], rd[1], rd[2], rd[3]};
    float kp_max = rd[4];
#pragma unroll
    for (int k = 1; k < K; ++k) kp_max = fmaxf(kp_max, rd[4 + k]);
    float ctip[4] = {r_ctip.x, r_ctip.y, r_ctip.z, r_ctip.w};
    const float coef[4] = {r_coef.x, r_coef.y, r_coef.z, r_coef.w};
    // lazy initialisation of the per-episode closest / furthest distances (:1044-1049, :1077-1079)
#pragma unroll
    for (int i = 0; i < 4; ++i) ctip[i] = (ctip[i] < 0.0f) ? cur[i] : ctip[i];
    float furthest = (r_furthest < 0.0f) ? cur[0] : r_furthest;
    float ckp = (r_ckp < 0.0f) ? kp_max : r_ckp;

    const long long progress = r_progress + (mode == kPostFull ? 1 : 0);   // progress_buf += 1 (:1429)
    // observation slots that see pre-update state (SURVEY 8a ordering notes)
    o[L.closest_kp] = ckp;
#pragma unroll
    for (int i = 0; i < 4; ++i) o[L.closest_tip + i] = ctip[i];
    o[L.lifted] = r_lifted ? 1.0f : 0.0f;
    o[L.progress] = logf((float)progress / 10.0f + 1.0f);
    o[L.successes] = logf(r_succ + 1.0f);

    if (mode == kPostObserveOnly) {
      o[L.reward] = r_old_rew_obs;
      if (valid) {
        b.closest_keypoint_max_dist[env] = ckp;
        b.furthest_hand_dist[env] = furthest;
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
      }
    } else {
      const float obj_z = s.obj[lane * 13 + 2];
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
      const bool near_goal = kp_max <= b.dynamic->keypoint_success_tolerance;
      const int near_steps = r_near + (near_goal ? 1 : 0);
      const bool is_success = near_steps >= p.success_steps;
      const float succ = r_succ + (is_success ? 1.0f : 0.0f);
      const float raw_ft = ft, raw_hp = hp, raw_lift = lifting_rew, raw_kp = kp_rew;
      ft *= p.distance_delta_rew_scale;
      hp *= p.hand_delta_rew_scale;
      lifting_rew *= p.lifting_rew_scale;
      kp_rew *= p.keypoint_rew_scale;
      const float kuka_pen = -1.0f * (rd[8] * p.kuka_actions_penalty_scale);
      const float allegro_pen = -1.0f * (rd[9] * p.allegro_actions_penalty_scale);
      const float bonus_rew = (near_goal ? 1.0f : 0.0f) * p.bonus_per_step;
      const float reward = ft + hp + lifting_rew + lift_bonus_rew + kp_rew + kuka_pen + allegro_pen + bonus_rew;
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
      const float true_obj = succ * b.dynamic->true_objective_scale + b.dynamic->true_objective_offset;
      const float rew_obs = reward * 0.01f;                          // (:1437-1439)
      o[L.reward] = rew_obs;

      float* tm = s.terms;
      tm[HA_KUKA_RAW_FINGERTIP_DELTA * kEnvsPerCta + lane] = raw_ft;
      tm[HA_KUKA_RAW_HAND_DELTA * kEnvsPerCta + lane] = raw_hp;
      tm[HA_KUKA_RAW_LIFTING * kEnvsPerCta + lane] = raw_lift;
      tm[HA_KUKA_RAW_KEYPOINT * kEnvsPerCta + lane] = raw_kp;
      tm[HA_KUKA_FINGERTIP_DELTA * kEnvsPerCta + lane] = ft;
      tm[HA_KUKA_HAND_DELTA * kEnvsPerCta + lane] = hp;
      tm[HA_KUKA_LIFTING * kEnvsPerCta + lane] = lifting_rew;
      tm[HA_KUKA_LIFT_BONUS * kEnvsPerCta + lane] = lift_bonus_rew;
      tm[HA_KUKA_KEYPOINT * kEnvsPerCta + lane] = kp_rew;
      tm[HA_KUKA_BONUS * kEnvsPerCta + lane] = bonus_rew;
      tm[HA_KUKA_KUKA_PENALTY * kEnvsPerCta + lane] = kuka_pen;
      tm[HA_KUKA_ALLEGRO_PENALTY * kEnvsPerCta + lane] = allegro_pen;

      const bool timeout_flag = timed_out && (resets != 0);          // vec_task.py:424
      if (valid) {
        b.rew_buf[env] = reward;
        b.reset_buf[env] = resets;
        b.reset_goal_buf[env] = is_success ? 1 : 0;
        b.progress_buf[env] = prog;
        if (b.randomize_buf) b.randomize_buf[env] += 1;              // (:1430)
        b.timeout_buf[env] = timeout_flag ? 1 : 0;
        b.successes[env] = succ;
        b.true_objective[env] = true_obj;
        b.near_goal_steps[env] = near_steps;
        b.lifted_object[env] = lifted ? 1 : 0;
        b.closest_keypoint_max_dist[env] = ckp;
        b.furthest_hand_dist[env] = furthest;
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
      }
      // per-CTA partial statistics, fixed order
      float st[kNPart];
      st[HA_KUKA_STAT_N] = valid ? 1.f : 0.f;
      st[HA_KUKA_STAT_SUM_REWARD] = valid ? reward : 0.f;
      st[HA_KUKA_STAT_SUM_RESETS] = valid ? (float)resets : 0.f;
      st[HA_KUKA_STAT_SUM_SUCCESS_EVENTS] = (valid && is_success) ? 1.f : 0.f;
      st[HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES] = valid ? r_prev_ep : 0.f;
      st[HA_KUKA_STAT_SUM_TRUE_OBJECTIVE] = valid ? true_obj : 0.f;
      st[HA_KUKA_STAT_MAX_TRUE_OBJECTIVE] = valid ? true_obj : -3.0e38f;
      st[HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE] = valid ? -true_obj : -3.0e38f;
      st[HA_KUKA_STAT_SUM_TIMEOUTS] = (valid && timeout_flag) ? 1.f : 0.f;
      st[HA_KUKA_STAT_SUM_LIFTED] = (valid && lifted) ? 1.f : 0.f;
#pragma unroll
      for (int i = 0; i < kNPart; ++i) {
        const bool is_max = (i == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE || i == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE);
        const float r = is_max ? warp_max(st[i]) : warp_sum(st[i]);
        if (lane == 0) b.scratch[(size_t)i * gridDim.x + blockIdx.x] = r;   // [stat][cta]
      }
    }
  }
  __syncthreads();

  // ------------------------------ phase S ---------------------------------
  if (mode == kPostFull && valid) {
    // rewards_episode[k] += term (:871-874, :923-926); episode_cumulative[k] = term (:925)
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int k = warp + 4 * j;
      const float t = s.terms[k * kEnvsPerCta + lane];
      b.rewards_episode[(size_t)k * n + env] = r_acc[j] + t;
    }
    // step terms in the reference's list order (:912-921)
    const int cum_src[HA_KUKA_N_STEP_TERMS] = {HA_KUKA_FINGERTIP_DELTA, HA_KUKA_HAND_DELTA, HA_KUKA_LIFTING, HA_KUKA_LIFT_BONUS,
                                               HA_KUKA_KEYPOINT, HA_KUKA_KUKA_PENALTY, HA_KUKA_ALLEGRO_PENALTY, HA_KUKA_BONUS};
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = warp + 4 * j;
      b.episode_cumulative[(size_t)k * n + env] = s.terms[cum_src[k] * kEnvsPerCta + lane];
    }
  }
  {
    // obs tile: clamp (:1174-1176 clamp_obs, vec_task.py:432 clip) and store contiguously
    const float c = p.clamp_abs_obs;
    float* dst = b.obs_buf + (size_t)env0 * OBS;
    const int total = nvalid * OBS;
    const int nvec = ((reinterpret_cast<uintptr_t>(dst) & 15) == 0) ? (total >> 2) : 0;
    for (int q = tid; q < nvec; q += kThreads) {
      float4 v = *reinterpret_cast<const float4*>(s.obs + 4 * q);
      if (c > 0.0f) {
        v.x = fminf(fmaxf(v.x, -c), c); v.y = fminf(fmaxf(v.y, -c), c);
        v.z = fminf(fmaxf(v.z, -c), c); v.w = fminf(fmaxf(v.w, -c), c);
      }
      *reinterpret_cast<float4*>(dst + 4 * q) = v;
    }
    for (int f = (nvec << 2) + tid; f < total; f += kThreads) {
      float v = s.obs[f];
      if (c > 0.0f) v = fminf(fmaxf(v, -c), c);
      dst[f] = v;
    }
  }

  // ---------------------- last CTA: final statistics ----------------------
  if (mode != kPostFull) return;
  __threadfence();
  if (tid == 0) {
    const unsigned t = atomicAdd(b.tickets + 0, 1u);
    s.last_flag = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (!s.last_flag) return;
  __threadfence();
  {
    // 128 threads: stat = tid / 8 (16 slots), 8 lanes per stat stride over the CTAs
    const int stat = tid >> 3, sub = tid & 7;
    const bool is_max = (stat == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE || stat == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE);
    float acc = is_max ? -3.0e38f : 0.f;
    if (stat < kNPart) {
      const volatile float* src = b.scratch + (size_t)stat * gridDim.x;
      for (int i = sub; i < (int)gridDim.x; i += 8) acc = is_max ? fmaxf(acc, src[i]) : acc + src[i];
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) {
      const float other = __shfl_xor_sync(0xffffffffu, acc, o);
      acc = is_max ? fmaxf(acc, other) : acc + other;
    }
    if (sub == 0 && stat < kNPart) { b.stats[stat] = acc; s.part[stat] = acc; }
    __syncthreads();
    if (tid == 0) {
      // means / min for extras (allegro_kuka_base.py:903-910), so the host needs no extra launches
      const float inv_n = 1.0f / s.part[HA_KUKA_STAT_N];
      b.stats[HA_KUKA_STAT_MEAN_REWARD] = s.part[HA_KUKA_STAT_SUM_REWARD] * inv_n;
      b.stats[HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES] = s.part[HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES] * inv_n;
      b.stats[HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE] = s.part[HA_KUKA_STAT_SUM_TRUE_OBJECTIVE] * inv_n;
      b.stats[HA_KUKA_STAT_MIN_TRUE_OBJECTIVE] = -s.part[HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE];
      b.tickets[0] = 0;                    // re-arm for the next launch (graph replay safe)
      // the reset/index counters of the pre-step are consumed by now: clear for the next pre-step
      b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    }
  }
}

}  // namespace ha
