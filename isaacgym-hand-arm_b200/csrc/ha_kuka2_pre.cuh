// K_pre for the two-arm AllegroKuka regrasping task: one launch replaces allegro_kuka_two_arms.py:1298-1351
// (pre_physics_step) including reset_target_pose (:1152-1157) with the regrasping _reset_target
// (allegro_kuka_two_arms_regrasping.py:75-110: +-0.75 m goal side, the object is put back on the table as well),
// reset_idx (:1215-1296), reset_object_pose (:1159-1193: +-0.35 m object side), the compacted actor rows for the
// *_indexed setters (object + goal ball rows; two arm rows per reset env), the all-absolute action -> DOF-target path
// (:1313-1334) and the random object forces (:1338-1351).
//
// Mapping as in ha_kuka_pre.cuh: CTA = 32-env tile, 128 threads; all threads walk the contiguous [32, 46] tiles of
// actions / prev_targets / cur_targets as 128-bit vectors; warp 0 does the per-env scalar reset work (lane = env);
// reset envs get their 46 DOFs re-randomised by one warp per env through shared memory; warp 1 applies the forces.
#pragma once
#include "ha_device.cuh"
#include "ha_kuka_pre.cuh"

namespace ha {

struct Kuka2PreSmem {
  float pos[kEnvsPerCta][48];
  long long fold[kThreads / 32];      // statistics fold of the CTAs below kNPart (kuka_fold_*)
};

__global__ void __launch_bounds__(kThreads)
kuka2_pre_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const float* __restrict__ actions,
                 const float* __restrict__ ext_rand, const int mode) {
  constexpr int D = 46;
  constexpr int kVecPerTile = kEnvsPerCta * D / 4;                  // 368
  constexpr int kIters = (kVecPerTile + kThreads - 1) / kThreads;   // 3
  __shared__ Kuka2PreSmem s;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.n_envs;
  const int env0 = blockIdx.x * kEnvsPerCta;
  const int nvalid = min(kEnvsPerCta, n - env0);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;
  const bool reset_only = (mode == kPreResetOnly);
  const size_t base = (size_t)env0 * D;
  const int total = nvalid * D, nvec = total >> 2;

  // CTAs 0 .. kNPart-1 also fold one statistic of the previous post step each (deferred statistics, see kuka_pre_kernel)
  const bool fold = blockIdx.x < kNPart && (n + kTile - 1) / kTile <= HA_KUKA_STAT_SLOTS;      // uniform over the CTA
  KukaStatFold sf;
  if (fold) kuka_fold_load(b, blockIdx.x, tid, sf);
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
  KukaRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_KUKA2_RND_STRIDE(D);
  R.rng.seed = b.dynamic->rng_seed;
  R.rng.step = b.dynamic->rng_step;
  const bool rst = valid && b.reset_buf[env] != 0;
  const bool grst = valid && !reset_only && b.reset_goal_buf[env] != 0;
  const unsigned m_rst = __ballot_sync(0xffffffffu, rst);
  const unsigned m_goal = __ballot_sync(0xffffffffu, grst);
  const unsigned m_any = m_rst | m_goal;

  // ---------------- DOF re-randomisation of reset envs (:1238-1258): one warp per env, lanes stride the 46 DOFs ----------------
  if (m_rst != 0u) {
    unsigned m = m_rst;
    int rank = 0;
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1;
      if ((rank++ & (kThreads / 32 - 1)) != warp) continue;
      const int re = env0 + e;
      for (int d = lane; d < D; d += 32) {
        const float u_pos = R.u(re, HA_KUKA2_RND_DOF_POS + d), u_vel = R.u(re, HA_KUKA2_RND_DOF_POS + D + d);
        const float lo = p.dof_lower[d], hi = p.dof_upper[d], dflt = p.dof_default[d];
        const float delta_max = hi - dflt, delta_min = lo - dflt;
        const float rand_delta = delta_min + (delta_max - delta_min) * u_pos;
        const float pos = dflt + p.dof_reset_noise[d] * rand_delta;
        const float vel = p.reset_dof_vel_noise * (2.0f * u_vel + -1.0f);
        const size_t idx = (size_t)re * D + d;
        *reinterpret_cast<float2*>(b.dof_state + 2 * idx) = make_float2(pos, vel);
        s.pos[e][d] = pos;
        if (reset_only) { b.prev_targets[idx] = pos; b.cur_targets[idx] = pos; }
      }
    }
  }
  if (fold) kuka_fold_warp(sf, blockIdx.x, tid, s.fold);
  __syncthreads();
  if (fold && tid == kThreads - 1) kuka_fold_finish(b, blockIdx.x, n, s.fold);

  // ---------------- all threads: every DOF is position-controlled (scale, moving average, clamp) ----------------
  if (!reset_only) {
    const float ema = p.act_moving_average, one_m_ema = p.one_minus_act_moving_average;
    const bool copy_actions = (actions != b.actions);
    auto target = [&](float av, float pv, int d) {
      const float lo = p.dof_lower[d], hi = p.dof_upper[d];
      float x = scale_to(av, lo, hi);
      x = ema * x + one_m_ema * pv;
      return clamp_to(x, lo, hi);
    };
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
        const float pv = ((m_rst >> e) & 1u) ? s.pos[e][d] : pvv[c];
        cur[c] = target(av[c], pv, d);
        if (++d == D) { d = 0; ++e; }
      }
      const float4 o = make_float4(cur[0], cur[1], cur[2], cur[3]);
      *(reinterpret_cast<float4*>(b.cur_targets + base) + q) = o;
      *(reinterpret_cast<float4*>(b.prev_targets + base) + q) = o;
      if (copy_actions) *(reinterpret_cast<float4*>(b.actions + base) + q) = a[i];
    }
    const int f = (nvec << 2) + tid;                  // ragged last tile: up to 3 trailing elements
    if (f < total) {
      const int e = f / D, d = f - e * D;
      const float av = __ldg(actions + base + f);
      const float pv = ((m_rst >> e) & 1u) ? s.pos[e][d] : b.prev_targets[base + f];
      const float x = target(av, pv, d);
      b.cur_targets[base + f] = x;
      b.prev_targets[base + f] = x;
      if (copy_actions) b.actions[base + f] = av;
    }
  }

  // ---------------- warp 1: random forces on the object body ----------------
  if (warp == 1 && valid) {
    float* frow = b.rb_forces + ((size_t)env * p.n_bodies + p.object_body) * 3;
    if (p.force_enabled && !reset_only) {
      float prob = b.random_force_prob[env];
      if (rst) prob = force_prob_from(p, R.u(env, HA_KUKA2_RND_FORCE_PROB));
      float f[3] = {0.f, 0.f, 0.f};
      if (!rst) { f[0] = frow[0]; f[1] = frow[1]; f[2] = frow[2]; }          // in flight under the generator pass
      const bool hit = R.ui(env, HA_KUKA2_RND_FORCE_TEST) < prob;
      float kick[3] = {0.f, 0.f, 0.f};
      if (hit) R.nrm3(env, HA_KUKA2_RND_FORCE_NORMAL, kick);                  // one counter block in fast mode
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        float v = f[c];                                                       // rb_forces[env_ids] = 0 on reset
        v *= p.force_decay_pow;
        if (hit) v = kick[c] * p.force_gain[0] * p.force_gain[1];
        frow[c] = v;
      }
    } else if (rst) {
      frow[0] = 0.f; frow[1] = 0.f; frow[2] = 0.f;
    }
  }
  // ---------------- warps 2,3: zero the per-episode accumulators of reset envs (:1291-1293) ----------------
  if (warp >= 2 && rst) {
    for (int k = warp - 2; k < HA_KUKA2_N_EPISODE_TERMS; k += 2) b.rewards_episode[(size_t)k * n + env] = 0.0f;
  }

  // ---------------- warp 0: per-env scalar reset work, lane = env ----------------
  if (warp != 0 || m_any == 0u) return;
  const int n_any = __popc(m_any), n_rst = __popc(m_rst);
  // reorientation (allegro_kuka_two_arms_reorientation.py:97-137): a new goal is a position + a rotation and leaves the object alone,
  // so a goal-only reset contributes one actor row; regrasping (allegro_kuka_two_arms_regrasping.py:75-110) also puts the object back
  const bool reorient = p.subtask == HA_KUKA_REORIENTATION;
  const unsigned m_goal_only = m_any & ~m_rst;
  const int rows_goal_only = reorient ? 1 : 2;
  int root_base = 0, hand_base = 0;
  if (lane == 0) {
    root_base = atomicAdd(b.reset_counts + 0, 2 * n_rst + rows_goal_only * __popc(m_goal_only));
    if (n_rst) { hand_base = atomicAdd(b.reset_counts + 1, 2 * n_rst); atomicAdd(b.reset_counts + 2, n_rst); }
    atomicAdd(b.reset_counts + 3, n_any);
  }
  root_base = __shfl_sync(0xffffffffu, root_base, 0);
  hand_base = __shfl_sync(0xffffffffu, hand_base, 0);
  if (!(rst || grst)) return;
  const int goal_row = p.root_regular ? env * p.actors_per_env + p.goal_slot : b.goal_actor[env];
  const int obj_row = p.root_regular ? env * p.actors_per_env + p.object_slot : b.object_actor[env];
  {
    // _reset_target: target position in the volume, shifted towards one of the arms
    const int g = rst ? HA_KUKA2_RND_RESET_GOAL : HA_KUKA2_RND_GOAL;
    float coords[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) coords[c] = p.target_min[c] + R.u(env, g + c) * p.target_size[c];
    const float side = 2.0f * R.u(env, g + 3) + -1.0f;
    coords[0] += (side > 0.0f) ? p.two_arm_x_ofs : -p.two_arm_x_ofs;
    float* gs = b.goal_states + (size_t)env * 13;
    float* rr = b.root_state + (size_t)goal_row * 13;
#pragma unroll
    for (int c = 0; c < 3; ++c) { gs[c] = coords[c]; rr[c] = coords[c]; }
    if (reorient) {                                        // get_random_quat (allegro_kuka_two_arms.py:1139-1150), goal row velocities = 0
      const f4 gq = random_quat(R.u(env, g + 4), R.u(env, g + 5), R.u(env, g + 6));
      gs[3] = gq.x; gs[4] = gq.y; gs[5] = gq.z; gs[6] = gq.w;
      rr[3] = gq.x; rr[4] = gq.y; rr[5] = gq.z; rr[6] = gq.w;
#pragma unroll
      for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
    }
    if (!reorient || rst) {
    // reset_object_pose: the draws of the last call win (goal block for a goal-only reset, the reset_idx block otherwise)
    const int ob = rst ? HA_KUKA2_RND_RESET_OBJ : g + 4;
    const float oside = 2.0f * R.u(env, ob) + -1.0f;
    const float* init = b.object_init_state + (size_t)env * 13;
    float* orow = b.root_state + (size_t)obj_row * 13;
    const float x_pos = (oside > 0.0f) ? p.two_arm_obj_x_ofs : -p.two_arm_obj_x_ofs;
    orow[0] = x_pos + p.reset_pos_noise[0] * (2.0f * R.u(env, ob + 1) + -1.0f);
    orow[1] = __ldg(init + 1) + p.reset_pos_noise[1] * (2.0f * R.u(env, ob + 2) + -1.0f);
    orow[2] = __ldg(init + 2) + p.reset_pos_noise[2] * (2.0f * R.u(env, ob + 3) + -1.0f);
    const f4 q = random_quat(R.u(env, ob + 4), R.u(env, ob + 5), R.u(env, ob + 6));
    orow[3] = q.x; orow[4] = q.y; orow[5] = q.z; orow[6] = q.w;
#pragma unroll
    for (int c = 7; c < 13; ++c) orow[c] = 0.0f;
    *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 8) = make_float4(-1.f, -1.f, -1.f, -1.f);
    *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 8 + 4) = make_float4(-1.f, -1.f, -1.f, -1.f);
    b.lifted_object[env] = 0;
    }
    b.reset_goal_buf[env] = 0;
    b.near_goal_steps[env] = 0;
    b.closest_keypoint_max_dist[env] = -1.0f;
  }
  const unsigned below = (1u << lane) - 1u;
  int k = root_base + 2 * __popc(m_rst & below) + rows_goal_only * __popc(m_goal_only & below);
  if (rst || !reorient) b.root_reset_indices[k++] = obj_row;
  b.root_reset_indices[k] = goal_row;
  if (rst) {
    b.random_force_prob[env] = force_prob_from(p, R.u(env, HA_KUKA2_RND_FORCE_PROB));
    b.progress_buf[env] = 0;
    b.reset_buf[env] = 0;
    b.prev_episode_successes[env] = b.successes[env];
    b.successes[env] = 0.0f;
    b.prev_episode_true_objective[env] = b.true_objective[env];
    b.true_objective[env] = 0.0f;
    const int h = hand_base + 2 * __popc(m_rst & below);
    b.hand_reset_indices[h] = b.hand_actor[2 * env];
    b.hand_reset_indices[h + 1] = b.hand_actor[2 * env + 1];
  }
}

}  // namespace ha
