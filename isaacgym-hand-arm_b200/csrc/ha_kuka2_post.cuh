// K_post for the two-arm AllegroKuka tasks (regrasping K = 1, reorientation K = 4): one launch replaces
// allegro_kuka_two_arms.py:1353-1370 (post_physics_step) = compute_observations (:939-1039) + compute_full_state
// (:1041-1133) + compute_kuka_reward (:818-878; _lifting_reward :763-787, _distance_delta_rewards :743-761 -- summed over
// both arms, no finger coefficients, no lifted gate -- _keypoint_reward :789-803, _compute_resets :805-813) + clamp_obs,
// plus the VecTask.step tail (vec_task.py:424-432).
//
// Same mapping as the single-arm kernel (ha_kuka_post.cuh): 32-env tile per CTA, cp.async staging of the AoS rows
// (2 palm rows, object root row, 8 fingertip rows, goal pose) and of the contiguous tiles (92-float DOF rows, keypoint
// table, scales), lane = env role warps (fingertips of both arms / keypoints / DOF tile / state + reward), the
// 32 x obs_dim tile leaves as one contiguous run of 128-bit stores, statistics through order-independent integer atomics.
#pragma once
#include "ha_device.cuh"
#include "ha_kuka_post.cuh"

namespace ha {

constexpr int kA2 = 2;                   // arms
constexpr int kD2 = 46;                  // DOFs
constexpr int kT2 = 8;                   // fingertips

struct Kuka2Layout { int dof_pos, dof_vel, palm_pos, palm_rot_vel, obj_rot_vel, tip_rel_palm, kp_rel_palm, kp_rel_goal, scales, closest_kp, lifted, progress, successes, reward, dim; };
__host__ __device__ constexpr Kuka2Layout kuka2_layout(int K) {   // allegro_kuka_two_arms.py:1041-1133
  Kuka2Layout l{};
  int o = 0;
  l.dof_pos = o; o += kD2;
  l.dof_vel = o; o += kD2;
  l.palm_pos = o; o += 3 * kA2;
  l.palm_rot_vel = o; o += 10 * kA2;
  l.obj_rot_vel = o; o += 10;
  l.tip_rel_palm = o; o += 3 * kT2;
  l.kp_rel_palm = o; o += 3 * kA2 * K;
  l.kp_rel_goal = o; o += 3 * K;
  l.scales = o; o += 3;
  l.closest_kp = o; o += 1;
  l.lifted = o; o += 1;
  l.progress = o; o += 1;
  l.successes = o; o += 1;
  l.reward = o; o += 1;
  l.dim = o;
  return l;
}

template <int K>
struct Kuka2PostSmem {
  static constexpr int kObs = kuka2_layout(K).dim;
  static constexpr int kPalmW = 13 * kA2 + 1, kTipW = 7 * kT2 + 1;      // odd row strides
  float dof[kTile * 2 * kD2];            // contiguous images first (16-byte copies)
  float kp[kTile * 3 * K + (K == 1 ? 32 : 0)];
  float scales[kTile * 3];
  float palm[kTile * kPalmW];            // [e][a * 13 + c]
  float obj[kTile * 13];
  float tip[kTile * kTipW];              // [e][t * 7 + c]
  float goal[kTile * 7];
  float xch[(kT2 + 1) * kTile];          // [q][e]: cur[0..7], kp_max
  float obs[kTile * kObs + 3];
};

template <int K>
__global__ void __launch_bounds__(kPostThreads)
kuka2_post_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const int mode) {
  using SM = Kuka2PostSmem<K>;
  constexpr int OBS = SM::kObs, PW = SM::kPalmW, TW = SM::kTipW;
  constexpr Kuka2Layout L = kuka2_layout(K);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SM& S = *reinterpret_cast<SM*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = ((tid >> 5) + blockIdx.x) & 3;   // roles rotate with the CTA index
  const int n = p.n_envs;
  const int env0 = blockIdx.x * kTile;
  const int nvalid = min(kTile, n - env0), last = nvalid - 1;
  const bool valid = lane < nvalid;
  const int env = env0 + min(lane, last);

  // ---------------- load ----------------
  {
    const size_t rbs = (size_t)p.n_bodies * 13;
    const float* rb0 = b.rigid_body_state + (size_t)env0 * rbs;
    for (int j = tid; j < kTile * 13 * kA2; j += kPostThreads) {
      const int e = j / (13 * kA2), r = j - 13 * kA2 * e, a = r / 13, c = r - 13 * a;
      cp_async4(S.palm + e * PW + r, rb0 + (size_t)min(e, last) * rbs + (size_t)p.palm_body[a] * 13 + c);
    }
    for (int j = tid; j < kTile * 7 * kT2; j += kPostThreads) {
      const int e = j / (7 * kT2), r = j - 7 * kT2 * e, t = r / 7, c = r - 7 * t;
      cp_async4(S.tip + e * TW + r, rb0 + (size_t)min(e, last) * rbs + (size_t)p.tip_body[t] * 13 + c);
    }
    for (int j = tid; j < kTile * 13; j += kPostThreads) {
      const int e = j / 13, c = j - 13 * e, ev = env0 + min(e, last);
      const size_t row = p.root_regular ? (size_t)ev * p.actors_per_env + p.object_slot : (size_t)__ldg(b.object_actor + ev);
      cp_async4(S.obj + j, b.root_state + row * 13 + c);
    }
    for (int j = tid; j < kTile * 7; j += kPostThreads) {
      const int e = j / 7, c = j - 7 * e;
      cp_async4(S.goal + j, b.goal_states + (size_t)(env0 + min(e, last)) * 13 + c);
    }
    const float* kp_src = b.object_keypoint_offsets + (size_t)env0 * 3 * K;
    const float* sc_src = b.object_scales + (size_t)env0 * 3;
    const float* dof_src = b.dof_state + (size_t)env0 * (2 * kD2);
    if (nvalid == kTile) {
      for (int q = tid; q < kTile * 3 * K / 4; q += kPostThreads) cp_async16(reinterpret_cast<float4*>(S.kp) + q, reinterpret_cast<const float4*>(kp_src) + q);
      for (int q = tid; q < kTile * 3 / 4; q += kPostThreads) cp_async16(reinterpret_cast<float4*>(S.scales) + q, reinterpret_cast<const float4*>(sc_src) + q);
      for (int q = tid; q < kTile * 2 * kD2 / 4; q += kPostThreads) cp_async16(reinterpret_cast<float4*>(S.dof) + q, reinterpret_cast<const float4*>(dof_src) + q);
    } else {
      for (int j = tid; j < nvalid * 3 * K; j += kPostThreads) cp_async4(S.kp + j, kp_src + j);
      for (int j = tid; j < nvalid * 3; j += kPostThreads) cp_async4(S.scales + j, sc_src + j);
      for (int j = tid; j < nvalid * 2 * kD2; j += kPostThreads) cp_async4(S.dof + j, dof_src + j);
    }
    cp_async_commit();
  }
  // per-env state of the state warp (lane = env)
  long long r_progress = 0, r_reset = 0, r_randomize = 0;
  float r_succ = 0.f, r_ckp = 0.f, r_initz = 0.f, r_prev_ep = 0.f, r_old_rew = 0.f, r_acc[HA_KUKA2_N_EPISODE_TERMS];
  float4 r_ct0 = make_float4(0.f, 0.f, 0.f, 0.f), r_ct1 = r_ct0;
  int r_near = 0;
  unsigned char r_lifted = 0;
  float keypoint_tol = 0.f, tobj_scale = 0.f, tobj_offset = 0.f;
  if (warp == 3) {
    keypoint_tol = b.dynamic->keypoint_success_tolerance;
    tobj_scale = b.dynamic->true_objective_scale;
    tobj_offset = b.dynamic->true_objective_offset;
    r_progress = b.progress_buf[env]; r_reset = b.reset_buf[env]; r_succ = b.successes[env];
    r_ckp = b.closest_keypoint_max_dist[env]; r_near = b.near_goal_steps[env]; r_lifted = b.lifted_object[env];
    r_initz = b.object_init_z ? ldg(b.object_init_z + env) : ldg(b.object_init_state + (size_t)env * 13 + 2);
    r_prev_ep = b.prev_episode_successes[env];
    r_ct0 = *reinterpret_cast<const float4*>(b.closest_fingertip_dist + (size_t)env * kT2);
    r_ct1 = *reinterpret_cast<const float4*>(b.closest_fingertip_dist + (size_t)env * kT2 + 4);
    if (mode == kPostObserveOnly) {
      r_old_rew = b.obs_buf[(size_t)env * OBS + L.reward];
    } else {
#pragma unroll
      for (int k = 0; k < HA_KUKA2_N_EPISODE_TERMS; ++k) r_acc[k] = b.rewards_episode[(size_t)k * n + env];
      if (b.randomize_buf) r_randomize = b.randomize_buf[env];
    }
  }
  cp_async_wait<0>();
  __syncthreads();

  float* o = S.obs + lane * OBS;
  const f3 palm_off = {p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]};
  float obj_z = 0.f;
  if (warp == 0) {
    // palm centres (:955-961) and fingertips of both arms (:971-996)
    const float* ob = S.obj + lane * 13;
    const f3 obj_pos = {ob[0], ob[1], ob[2]};
#pragma unroll
    for (int a = 0; a < kA2; ++a) {
      const float* pr = S.palm + lane * PW + a * 13;
      const f3 pc = f3{pr[0], pr[1], pr[2]} + quat_rotate(f4{pr[3], pr[4], pr[5], pr[6]}, palm_off);
      o[L.palm_pos + 3 * a + 0] = pc.x; o[L.palm_pos + 3 * a + 1] = pc.y; o[L.palm_pos + 3 * a + 2] = pc.z;
#pragma unroll
      for (int i = 0; i < kT1; ++i) {
        const int t = a * kT1 + i;
        const float* tr = S.tip + lane * TW + t * 7;
        const f3 to = f3{tr[0], tr[1], tr[2]} + quat_rotate(f4{tr[3], tr[4], tr[5], tr[6]}, f3{p.tip_offset[i][0], p.tip_offset[i][1], p.tip_offset[i][2]});
        const f3 rel_obj = to - obj_pos, rel_palm = to - pc;
        S.xch[t * kTile + lane] = norm3(rel_obj);
        o[L.tip_rel_palm + 3 * t + 0] = rel_palm.x; o[L.tip_rel_palm + 3 * t + 1] = rel_palm.y; o[L.tip_rel_palm + 3 * t + 2] = rel_palm.z;
      }
    }
  } else if (warp == 1) {
    // keypoints (:1003-1025): relative to the goal and to both palm centres ([arm][keypoint] order)
    const float* ob = S.obj + lane * 13;
    const float* gl = S.goal + lane * 7;
    const f3 obj_pos = {ob[0], ob[1], ob[2]}, goal_pos = {gl[0], gl[1], gl[2]};
    const f4 obj_rot = {ob[3], ob[4], ob[5], ob[6]}, goal_rot = {gl[3], gl[4], gl[5], gl[6]};
    f3 pc[kA2];
#pragma unroll
    for (int a = 0; a < kA2; ++a) {
      const float* pr = S.palm + lane * PW + a * 13;
      pc[a] = f3{pr[0], pr[1], pr[2]} + quat_rotate(f4{pr[3], pr[4], pr[5], pr[6]}, palm_off);
    }
    float kp_max = 0.f;
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const float* kq = S.kp + lane * 3 * K + 3 * k;
      const f3 off = {kq[0], kq[1], kq[2]};
      const f3 okp = obj_pos + quat_rotate(obj_rot, off);
      const f3 gkp = goal_pos + quat_rotate(goal_rot, off);
      const f3 rel_goal = okp - gkp;
      const float dist = norm3(rel_goal);
      kp_max = (k == 0) ? dist : fmaxf(kp_max, dist);
      o[L.kp_rel_goal + 3 * k + 0] = rel_goal.x; o[L.kp_rel_goal + 3 * k + 1] = rel_goal.y; o[L.kp_rel_goal + 3 * k + 2] = rel_goal.z;
#pragma unroll
      for (int a = 0; a < kA2; ++a) {
        const f3 rel_palm = okp - pc[a];
        const int s = L.kp_rel_palm + 3 * (a * K + k);
        o[s + 0] = rel_palm.x; o[s + 1] = rel_palm.y; o[s + 2] = rel_palm.z;
      }
    }
    S.xch[kT2 * kTile + lane] = kp_max;
  } else if (warp == 2) {
    const float2* dl = reinterpret_cast<const float2*>(S.dof) + lane * kD2;
#pragma unroll 2
    for (int d = 0; d < kD2; ++d) {
      const float2 pv = dl[d];
      const float lo = p.dof_lower[d], hi = p.dof_upper[d];
      o[L.dof_pos + d] = div_by_const(2.0f * pv.x - hi - lo, hi - lo, p.dof_range_rcp[d]);
      o[L.dof_vel + d] = pv.y;
    }
  } else {
    const float* ob = S.obj + lane * 13;
    obj_z = ob[2];
#pragma unroll
    for (int a = 0; a < kA2; ++a) {
      const float* pr = S.palm + lane * PW + a * 13;
#pragma unroll
      for (int c = 0; c < 10; ++c) o[L.palm_rot_vel + 10 * a + c] = pr[3 + c];
    }
#pragma unroll
    for (int c = 0; c < 10; ++c) o[L.obj_rot_vel + c] = ob[3 + c];
#pragma unroll
    for (int c = 0; c < 3; ++c) o[L.scales + c] = S.scales[lane * 3 + c];
  }
  __syncthreads();

  KukaPostAcc acc;
  if (warp == 3) {
    float cur[kT2], ctip[kT2] = {r_ct0.x, r_ct0.y, r_ct0.z, r_ct0.w, r_ct1.x, r_ct1.y, r_ct1.z, r_ct1.w};
#pragma unroll
    for (int t = 0; t < kT2; ++t) cur[t] = S.xch[t * kTile + lane];
    const float kp_max = S.xch[kT2 * kTile + lane];
#pragma unroll
    for (int t = 0; t < kT2; ++t) ctip[t] = (ctip[t] < 0.0f) ? cur[t] : ctip[t];      // lazy init (:982-984)
    float ckp = (r_ckp < 0.0f) ? kp_max : r_ckp;                                       // (:1029-1031)
    const long long progress = r_progress + (mode == kPostFull ? 1 : 0);
    o[L.closest_kp] = ckp;
    o[L.lifted] = r_lifted ? 1.0f : 0.0f;
    o[L.progress] = logf((float)progress / 10.0f + 1.0f);
    o[L.successes] = logf(r_succ + 1.0f);
    if (mode == kPostObserveOnly) {
      o[L.reward] = r_old_rew;
      if (valid) {
        b.closest_keypoint_max_dist[env] = ckp;
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * kT2) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * kT2 + 4) = make_float4(ctip[4], ctip[5], ctip[6], ctip[7]);
      }
    } else {
      const float z_lift = 0.05f + obj_z - r_initz;
      float lifting_rew = clipf(z_lift, 0.0f, 0.5f);
      const bool lifted_prev = r_lifted != 0;
      const bool lifted = (z_lift > p.lifting_bonus_threshold) || lifted_prev;
      const float lift_bonus_rew = p.lifting_bonus * ((lifted && !lifted_prev) ? 1.0f : 0.0f);
      lifting_rew *= lifted ? 0.0f : 1.0f;
      float ft = 0.f;
#pragma unroll
      for (int a = 0; a < kA2; ++a) {                 // sum over fingertips, then over arms (:752-753)
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < kT1; ++i) {
          const int t = a * kT1 + i;
          s += clipf(ctip[t] - cur[t], 0.0f, 10.0f);
          ctip[t] = fminf(ctip[t], cur[t]);
        }
        ft = (a == 0) ? s : ft + s;
      }
      float kp_rew = clipf(ckp - kp_max, 0.0f, 100.0f);
      ckp = fminf(ckp, kp_max);
      kp_rew *= lifted ? 1.0f : 0.0f;
      const bool near_goal = kp_max <= keypoint_tol;
      const int near_steps = r_near + (near_goal ? 1 : 0);
      const bool is_success = near_steps >= p.success_steps;
      const float succ = r_succ + (is_success ? 1.0f : 0.0f);
      float t8[HA_KUKA2_N_EPISODE_TERMS];
      t8[0] = ft; t8[1] = lifting_rew; t8[2] = kp_rew;
      ft *= p.distance_delta_rew_scale;
      lifting_rew *= p.lifting_rew_scale;
      kp_rew *= p.keypoint_rew_scale;
      const float bonus_rew = (near_goal ? 1.0f : 0.0f) * p.bonus_per_step;
      const float reward = ft + lifting_rew + lift_bonus_rew + kp_rew + bonus_rew;
      t8[3] = ft; t8[4] = lifting_rew; t8[5] = lift_bonus_rew; t8[6] = kp_rew; t8[7] = bonus_rew;
      long long resets = (obj_z < p.fall_z) ? 1 : r_reset;
      long long prog = progress;
      if (p.max_consecutive_successes > 0) {
        prog = is_success ? 0 : prog;
        resets = (succ >= (float)p.max_consecutive_successes) ? 1 : resets;
      }
      const bool timed_out = prog >= (long long)(p.max_episode_length - 1);
      resets = timed_out ? 1 : resets;
      const float true_obj = succ * tobj_scale + tobj_offset;
      const bool timeout_flag = timed_out && (resets != 0);
      o[L.reward] = reward * 0.01f;
      if (valid) {
        b.rew_buf[env] = reward;
        b.reset_buf[env] = resets;
        b.progress_buf[env] = prog;
        b.reset_goal_buf[env] = is_success ? 1 : 0;
        if (b.randomize_buf) b.randomize_buf[env] = r_randomize + 1;
        b.timeout_buf[env] = timeout_flag ? 1 : 0;
        b.successes[env] = succ;
        b.true_objective[env] = true_obj;
        b.near_goal_steps[env] = near_steps;
        b.lifted_object[env] = lifted ? 1 : 0;
        b.closest_keypoint_max_dist[env] = ckp;
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * kT2) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
        *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * kT2 + 4) = make_float4(ctip[4], ctip[5], ctip[6], ctip[7]);
#pragma unroll
        for (int k = 0; k < HA_KUKA2_N_EPISODE_TERMS; ++k) b.rewards_episode[(size_t)k * n + env] = r_acc[k] + t8[k];
#pragma unroll
        for (int k = 0; k < HA_KUKA2_N_STEP_TERMS; ++k) b.episode_cumulative[(size_t)k * n + env] = t8[3 + k];
        acc.rew = __float2ll_rn(reward * stat_scale(HA_KUKA_STAT_SUM_REWARD));
        acc.prev_ep = __float2ll_rn(r_prev_ep * stat_scale(HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES));
        acc.tobj = __float2ll_rn(true_obj * stat_scale(HA_KUKA_STAT_SUM_TRUE_OBJECTIVE));
        acc.mx = true_obj; acc.neg_mn = -true_obj;
        acc.n = 1; acc.resets = (resets != 0); acc.success = is_success; acc.timeout = timeout_flag; acc.lifted = lifted;
      }
    }
  }
  __syncthreads();
  // statistics: deferred like the single-arm kernel (one slot of partials per CTA, folded by the next pre step) whenever the
  // grid fits the slots; integer atomics + last-CTA ticket inside the launch otherwise
  const bool deferred = gridDim.x <= HA_KUKA_STAT_SLOTS;
  if (warp == 3 && mode == kPostFull) {
    if (deferred) {
      kuka_post_store_partials(b, acc, lane);
      if (blockIdx.x == 0 && lane == 0) {               // step bookkeeping, see kuka_post_kernel
        b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
        b.dynamic->rng_step = b.dynamic->rng_step + 1;
      }
    } else {
      kuka_post_stats(b, acc, lane);
    }
  } else {
    const float c = p.clamp_abs_obs;
    const bool do_clamp = c > 0.0f;
    float* dst = b.obs_buf + (size_t)env0 * OBS;
    const int total = nvalid * OBS, nvec = total >> 2;
    const int sid = warp * 32 + lane, nst = (mode == kPostFull) ? kPostThreads - 32 : kPostThreads;
    for (int q = sid; q < nvec; q += nst) {
      float4 v = reinterpret_cast<const float4*>(S.obs)[q];
      if (do_clamp) {
        v.x = fminf(fmaxf(v.x, -c), c); v.y = fminf(fmaxf(v.y, -c), c);
        v.z = fminf(fmaxf(v.z, -c), c); v.w = fminf(fmaxf(v.w, -c), c);
      }
      reinterpret_cast<float4*>(dst)[q] = v;
    }
    const int f = (nvec << 2) + sid;
    if (f < total) {
      float v = S.obs[f];
      if (do_clamp) v = fminf(fmaxf(v, -c), c);
      dst[f] = v;
    }
  }
}

}  // namespace ha
