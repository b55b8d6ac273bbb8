// K_pre of AllegroHandDextremeADR: one launch replaces pre_physics_step (allegro_hand_dextreme.py:912-949) -- the ADR
// apply_reset_buf bookkeeping (a reset is applied one step after the env was marked), reset_target_pose (:709-733),
// reset_idx (:775-858), apply_actions (:955-1015) with the ADR action latency / delay / affine + white noise
// (:1430-1441), and apply_random_forces (:1017-1030).  Masks are consumed in place (no nonzero / unique / len).
//
// The reference shifts the whole [N, L, 16] action queue every step (:969).  Here the queue is a ring buffer stored
// slot-major [L, N, 16]: the step writes one coalesced [N,16] slot and gathers each env's delayed action from slot
// (head - latency[env]) mod L; only envs whose episode restarted (progress == 0) rewrite their L rows.
//
// Mapping: CTA = 32-env tile, 128 threads.  Reset envs: one warp per env re-randomises the DOFs (lane = DOF); warp 0:
// per-env scalar reset work (lane = env); all threads: the [32,16] action tile flat (4 elements per thread);
// warp 1: random forces.  The last CTA publishes the new queue head and the count of pending resets.
#pragma once
#include "ha_dex_common.cuh"

namespace ha {

struct DexPreSmem {
  float pos[kDexTile][kDexD];     // re-randomised DOF positions of the tile's reset envs
  float stdev[kDexTile];          // white action noise
  float alpha[kDexTile];
  int lat[kDexTile];
  unsigned m_delay;
};

__global__ void __launch_bounds__(kDexThreads)
dex_pre_kernel(const __grid_constant__ HaDexParams p, const __grid_constant__ HaDexBuffers b, const float* __restrict__ actions,
               const float* __restrict__ rna, const float* __restrict__ ext_rand) {
  __shared__ DexPreSmem s;
  constexpr int D = kDexD;
  const unsigned FULL = 0xffffffffu;
  const int Dp = p.dofs_per_env, L = p.queue_len;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.n_envs;
  const int env0 = blockIdx.x * kDexTile;
  const int nvalid = min(kDexTile, n - env0);
  const int env = env0 + lane;
  const bool valid = lane < nvalid;

  DexRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_DEX_RND_STRIDE;
  R.space = 0;
  R.rng.seed = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_SEED];
  R.rng.step = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_STEP];
  const int head_prev = (int)b.dynamic[HA_DEX_DYN_QUEUE_HEAD];
  const int head = (head_prev + 1 >= L) ? 0 : head_prev + 1;
  const bool all_reset = b.dynamic[HA_DEX_DYN_N_APPLY] == (long long)n;        // len(env_ids) == num_envs (:848)

  // every warp reads the tile's masks (coalesced int64) -> identical ballots in all warps
  const long long v_reset = valid ? b.reset_buf[env] : 0, v_apply = valid ? b.apply_reset_buf[env] : 0;
  const bool rst = v_apply != 0;                                               // env_ids_reset = apply_reset_buf.nonzero() (:931)
  const bool mark = (v_reset & ~v_apply) != 0;                                 // env_mask_randomize (:930)
  const bool grst = valid && b.reset_goal_buf[env] != 0;
  long long progress = valid ? b.progress_buf[env] : 1;
  if (rst) progress = all_reset ? min((long long)(R.u(env, HA_DEX_RND_PROGRESS) * (float)p.max_episode_length), (long long)p.max_episode_length - 1) : 0;
  const unsigned m_rst = __ballot_sync(FULL, rst);
  const unsigned m_goal = __ballot_sync(FULL, grst);
  const unsigned m_mark = __ballot_sync(FULL, mark);
  const unsigned m_fresh = __ballot_sync(FULL, valid && progress == 0);        // refreshed = progress_buf == 0 (:959)
  const unsigned m_any = m_rst | m_goal;

  // ---------------- DOF re-randomisation of reset envs: one warp per env, lane = DOF (:812-826) ----------------
  if (m_rst != 0u) {
    unsigned m = m_rst;
    int rank = 0;
    while (m) {
      const int e = __ffs(m) - 1;
      m &= m - 1;
      if ((rank++ & (kDexThreads / 32 - 1)) != warp || lane >= D) continue;
      const int d = lane, re = env0 + e;
      const float r_pos = 2.0f * R.u(re, HA_DEX_RND_RESET + 5 + d) + -1.0f;
      const float r_vel = 2.0f * R.u(re, HA_DEX_RND_RESET + 5 + D + d) + -1.0f;
      const float lo = p.dof_lower[d], hi = p.dof_upper[d];
      const float delta_max = hi - 0.0f, delta_min = lo - 0.0f;               // hand_dof_default_pos is the zero vector (:274)
      const float rand_delta = delta_min + (delta_max - delta_min) * ((r_pos + 1.0f) / 2.0f);
      const float pos = 0.0f + p.reset_dof_pos_noise * rand_delta;
      const float vel = 0.0f + p.reset_dof_vel_noise * r_vel;
      *reinterpret_cast<float2*>(b.dof_state + 2 * ((size_t)re * Dp + d)) = make_float2(pos, vel);
      s.pos[e][d] = pos;
      b.prev_targets[(size_t)re * D + d] = pos;
      b.prev_prev_targets[(size_t)re * D + d] = pos;
    }
  }
  if (warp == 2) {                                                            // per-env scalars of the action path
    float dly = 1.0f, prob = 0.0f;
    if (valid) {
      dly = R.u(env, HA_DEX_RND_ACTION_DELAY);
      prob = __ldg(b.adr_action_delay_prob + env);
      s.stdev[lane] = dex_stdev(__ldg(b.adr_affine_action_white + env));
      s.alpha[lane] = rna ? __ldg(b.adr_rna_alpha + env) : 0.0f;
      s.lat[lane] = (int)__ldg(b.action_latency + env);
    }
    const unsigned m = __ballot_sync(FULL, valid && dly < prob);               // action_delay_mask (:1432)
    if (lane == 0) s.m_delay = m;
  }
  __syncthreads();

  // ---------------- all threads: action queue, latency / delay / noise, DOF targets (:955-1015, :1430-1441) ----------------
  if (tid < nvalid * 4) {                                                      // thread = 4 consecutive DOFs of one env (128-bit accesses)
    const bool copy_actions = (actions != b.actions);
    const size_t slot4 = (size_t)n * D / 4;                                    // float4 elements per queue slot
    const int e = tid >> 2, d0 = (tid & 3) * 4;
    const size_t g4 = (size_t)env0 * (D / 4) + tid;                            // float4 index into an [N,16] buffer
    const bool is_rst = (m_rst >> e) & 1u, fresh = (m_fresh >> e) & 1u, delayed = (s.m_delay >> e) & 1u;
    auto ld4 = [](const float* base, size_t i4) { return __ldg(reinterpret_cast<const float4*>(base) + i4); };
    auto st4 = [](float* base, size_t i4, float4 v) { reinterpret_cast<float4*>(base)[i4] = v; };
    float4* queue = reinterpret_cast<float4*>(b.action_queue);
    const float4 act = ld4(actions, g4);
    float lo[4], hi[4], rcp[4], av[4], prev[4], out[4], dd[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) { lo[c] = p.dof_lower[d0 + c]; hi[c] = p.dof_upper[d0 + c]; rcp[c] = p.dof_range_rcp[d0 + c]; }
    float4 unscaled = make_float4(0.f, 0.f, 0.f, 0.f);
    if (fresh) {                                                               // queue and prev_actions restart from the current pose
      float u[4];
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const float dp = is_rst ? s.pos[e][d0 + c] : b.dof_state[2 * ((size_t)(env0 + e) * Dp + d0 + c)];
        u[c] = div_by_const(2.0f * dp - hi[c] - lo[c], hi[c] - lo[c], rcp[c]);
      }
      unscaled = make_float4(u[0], u[1], u[2], u[3]);
      for (int k = 0; k < L; ++k)
        if (k != head) queue[(size_t)k * slot4 + g4] = unscaled;
    }
    queue[(size_t)head * slot4 + g4] = act;                                    // prev_actions_queue[:, 0] = actions
    const int lat = s.lat[e];
    float4 a4;
    if (delayed) a4 = fresh ? unscaled : *(reinterpret_cast<const float4*>(b.prev_actions) + g4);
    else if (lat == 0) a4 = act;
    else if (fresh) a4 = unscaled;
    else { int k = head - lat; k += (k < 0) ? L : 0; a4 = queue[(size_t)k * slot4 + g4]; }
    const float4 wn = R.nrm4(env0 + e, HA_DEX_RND_ACTION_WHITE + d0);
    const float4 sc = ld4(b.affine_actions_scaling, g4), ad = ld4(b.affine_actions_additive, g4);
    const float sd = s.stdev[e];
    av[0] = sc.x * a4.x + ad.x + wn.x * sd; av[1] = sc.y * a4.y + ad.y + wn.y * sd;
    av[2] = sc.z * a4.z + ad.z + wn.z * sd; av[3] = sc.w * a4.w + ad.w + wn.w * sd;
    if (rna) {                                                                 // rna_alpha blend (:887-890)
      const float al = s.alpha[e];
      const float4 rp = ld4(rna, g4);
      const float r[4] = {rp.x, rp.y, rp.z, rp.w};
#pragma unroll
      for (int c = 0; c < 4; ++c) av[c] = al * div_by_const(2.0f * r[c] - hi[c] - lo[c], hi[c] - lo[c], rcp[c]) + (1.0f - al) * av[c];
    }
    if (is_rst) {
#pragma unroll
      for (int c = 0; c < 4; ++c) prev[c] = s.pos[e][d0 + c];
    } else {
      const float4 pv = *(reinterpret_cast<const float4*>(b.prev_targets) + g4);
      prev[0] = pv.x; prev[1] = pv.y; prev[2] = pv.z; prev[3] = pv.w;
    }
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      float x = scale_to(av[c], lo[c], hi[c]);
      if (p.use_capped_dof_control) x = prev[c] + fminf(fmaxf(x - prev[c], -p.max_dof_delta), p.max_dof_delta);
      x = p.act_moving_average * x + p.one_minus_act_moving_average * prev[c];
      out[c] = clamp_to(x, lo[c], hi[c]);
      dd[c] = out[c] - prev[c];
    }
    st4(b.cur_targets, g4, make_float4(out[0], out[1], out[2], out[3]));
    st4(b.dof_delta, g4, make_float4(dd[0], dd[1], dd[2], dd[3]));
    st4(b.prev_actions, g4, act);
    if (copy_actions) st4(b.actions, g4, act);
  }

  // ---------------- warp 1: random forces on the object body (:1017-1030) ----------------
  if (warp == 1) {
    float* frow = b.rb_forces + ((size_t)env * p.n_bodies + p.object_body) * 3;
    if (p.force_enabled) {
      if (valid) {
        float prob = b.random_force_prob[env];
        if (rst) prob = expf(p.force_log_c0 * R.u(env, HA_DEX_RND_FORCE_PROB) + p.force_log_c1);
        float f[3] = {0.f, 0.f, 0.f};                                          // rb_forces[env_ids] = 0 (:784)
        if (!rst) { f[0] = frow[0]; f[1] = frow[1]; f[2] = frow[2]; }          // in flight under the generator pass
        const bool hit = R.u(env, HA_DEX_RND_FORCE_TEST) < prob;
        float kick[3] = {0.f, 0.f, 0.f};
        if (hit) R.nrm3(env, HA_DEX_RND_FORCE_NORMAL, kick);                   // one counter block in fast mode
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
  if (warp != 0) return;

  // ---------------- warp 0: per-env scalar reset work, lane = env ----------------
  const unsigned below = (1u << lane) - 1u;
  const unsigned m_goal_only = m_any & ~m_rst;
  int root_base = 0, hand_base = 0;
  if (m_any != 0u && lane == 0) {
    root_base = atomicAdd(b.reset_counts + 0, 2 * __popc(m_rst) + __popc(m_goal_only));
    if (m_rst) { hand_base = atomicAdd(b.reset_counts + 1, __popc(m_rst)); atomicAdd(b.reset_counts + 2, __popc(m_rst)); }
  }
  if (m_mark != 0u && lane == 0) atomicAdd(b.reset_counts + 3, __popc(m_mark));
  root_base = __shfl_sync(FULL, root_base, 0);
  hand_base = __shfl_sync(FULL, hand_base, 0);
  if (rst || grst) {
    const int goal_row = p.root_regular ? env * p.actors_per_env + p.goal_slot : b.goal_actor[env];
    {
      // reset_target_pose (:709-733); an env in both sets is drawn twice by the reference and the reset_idx draw wins
      const int slot = rst ? HA_DEX_RND_RESET_GOAL : HA_DEX_RND_GOAL;
      const float r0 = 2.0f * R.u(env, slot + 0) + -1.0f, r1 = 2.0f * R.u(env, slot + 1) + -1.0f;
      const f4 q = dex_new_rot(p.apply_random_quat, r0, r1, R.u(env, slot + 4), R.u(env, slot + 5), R.u(env, slot + 6));
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
      b.best_rotation_dist[env] = -1.0f;
    }
    int k = root_base + 2 * __popc(m_rst & below) + __popc(m_goal_only & below);
    if (rst) {
      // reset_idx: object pose (:786-799)
      const int obj_row = p.root_regular ? env * p.actors_per_env + p.object_slot : b.object_actor[env];
      const float* init = b.object_init_state + (size_t)env * 13;
      float* rr = b.root_state + (size_t)obj_row * 13;
      float r[5];
#pragma unroll
      for (int c = 0; c < 5; ++c) r[c] = 2.0f * R.u(env, HA_DEX_RND_RESET + c) + -1.0f;
      rr[0] = __ldg(init + 0) + p.reset_position_noise * r[0];
      rr[1] = __ldg(init + 1) + p.reset_position_noise * r[1];
      rr[2] = __ldg(init + 2) + p.reset_position_noise_z * r[2];
      const int qs = HA_DEX_RND_RESET + 2 * D + 5;
      const f4 q = dex_new_rot(p.apply_random_quat, r[3], r[4], R.u(env, qs), R.u(env, qs + 1), R.u(env, qs + 2));
      rr[3] = q.x; rr[4] = q.y; rr[5] = q.z; rr[6] = q.w;
#pragma unroll
      for (int c = 7; c < 13; ++c) rr[c] = 0.0f;
      b.random_force_prob[env] = expf(p.force_log_c0 * R.u(env, HA_DEX_RND_FORCE_PROB) + p.force_log_c1);      // (:807-809)
      // wrist-relative latches from the object pose as last observed and the current palm pose (:838-846)
      const float* palm = b.rigid_body_state + ((size_t)env * p.n_bodies + p.palm_body) * 13;
      float pp[7], op[7];
#pragma unroll
      for (int c = 0; c < 7; ++c) { pp[c] = __ldg(palm + c); op[c] = b.object_pose[(size_t)env * 7 + c]; }
      const Mat3 Rp = dex_quat_to_matrix(f4{pp[3], pp[4], pp[5], pp[6]});
      const Pose7 rel = dex_pose_wrt_wrist(op, Rp, pp);
#pragma unroll
      for (int c = 0; c < 7; ++c) { b.obs_object_pose[(size_t)env * 7 + c] = rel.v[c]; b.obs_object_pose_freq[(size_t)env * 7 + c] = rel.v[c]; }
      b.progress_buf[env] = progress;
      b.reset_buf[env] = 0;
      b.successes[env] = 0.0f;
      b.best_rotation_dist[env] = -1.0f;
      b.hold_count_buf[env] = 0;
      b.root_reset_indices[k++] = obj_row;
      b.hand_reset_indices[hand_base + __popc(m_rst & below)] = b.hand_actor[env];
    }
    b.root_reset_indices[k] = goal_row;
  }
  if (valid && (rst || mark)) b.apply_reset_buf[env] = mark ? 1 : 0;           // (:939) / (:853)
  // ---------------- pending-reset count and queue head, published by the last CTA ----------------
  unsigned long long* dyn = reinterpret_cast<unsigned long long*>(b.dynamic);
  const int n_apply = __popc(__ballot_sync(FULL, valid && (mark || (v_apply != 0 && !rst))));
  int is_last = 0;
  if (lane == 0) {
    const unsigned long long old = atomicAdd(dyn + HA_DEX_DYN_APPLY_ACC, (unsigned long long)n_apply);
    const unsigned dep = (old == 0xffffffffffffffffULL) ? 1u : 0u;             // the ticket depends on the returned atomic
    is_last = (atomicAdd(b.tickets + 1, 1u + dep) == gridDim.x - 1);
    if (is_last) {
      __threadfence();
      dyn[HA_DEX_DYN_N_APPLY] = atomicExch(dyn + HA_DEX_DYN_APPLY_ACC, 0ULL);
      dyn[HA_DEX_DYN_QUEUE_HEAD] = (unsigned long long)head;
      b.tickets[1] = 0;
    }
  }
}

}  // namespace ha
