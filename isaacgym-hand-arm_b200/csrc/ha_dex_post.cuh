// K_post of AllegroHandDextremeADR: one launch replaces post_physics_step (allegro_hand_dextreme.py:1033-1048) =
// compute_observations (base :606-692 with compute_poses_wrt_wrist :533-588, ADR extension :1443-1489 with
// get_random_cube_observation :739-771 and sample_gaussian_adr :1383-1387) + compute_reward / compute_hand_reward
// (:438-460, :1598-1663) + prev_targets update (:1046) + the dict-observation step tail (adr_vec_task.py:289-325:
// timeout_buf, clamp of all 20 entries).  The 20 dict entries are the column bands of ONE [N, 285] row buffer.
//
// Mapping: CTA = 32-env tile, 128 threads.
//   load      cp.async into shared memory: 16-byte copies of the contiguous per-env tiles (DOF state, DOF forces, force
//             sensors, actions, targets, the six affine tensors, randomisation parameters, the two pose latches), 4-byte
//             copies of the 52-byte AoS rows (object root row, goal pose, palm pose, fingertip rows, object force row);
//   compute   warp 0, lane = env: wrist-relative poses (closed-form rigid inverse), relative rotations, the ADR pose
//             pipeline (affine + white noise, random pose injection, refresh-rate and delay latches);
//             warp 1, lane = env: rotation / position distances, per-env reward state;
//             warps 2-3: four lanes per env, four DOFs per lane: randomised DOF positions and the three 16-term sums of the reward;
//   reward    warp 1 finishes compute_hand_reward and the statistics while
//   assemble  warps 0, 2, 3 walk the 32 x 285 tile flat through a column-descriptor table, apply scale / unscale / clip
//             and write 128-bit coalesced stores; the latches and prev_targets leave through the same flat path.
#pragma once
#include "ha_dex_common.cuh"

namespace ha {

enum { kDexPostFull = 0, kDexPostObserveOnly = 1 };
constexpr int kDexCmp = 51;        // computed floats per env (odd stride: lane = env writes are conflict-free)
enum { kCmpObjPose = 0, kCmpGoalPose = 7, kCmpGoalRel = 14, kCmpCam = 18, kCmpCamRel = 25, kCmpRotDist = 29, kCmpDofRand = 31, kCmpDelay = 47 };

struct DexStageOfs {
  int dof, frc, sens, act, cur, prev, a_as, a_aa, a_ps, a_pa, a_ds, a_da, cube, handp, grav, freq, cam, obj, goal, palm, tip, frow, cmp, sums, total;
};
__host__ __device__ inline DexStageOfs dex_stage_ofs() {
  DexStageOfs o{};
  int k = 0;
  auto take = [&](int per_env) { const int at = k; k += kDexTile * per_env; k = (k + 3) & ~3; return at; };
  o.dof = take(32); o.frc = take(16); o.sens = take(24); o.act = take(16); o.cur = take(16); o.prev = take(16);
  o.a_as = take(16); o.a_aa = take(16); o.a_ps = take(7); o.a_pa = take(7); o.a_ds = take(16); o.a_da = take(16);
  o.cube = take(3); o.handp = take(1); o.grav = take(3); o.freq = take(7); o.cam = take(7);
  o.obj = take(13); o.goal = take(7); o.palm = take(7); o.tip = take(52); o.frow = take(3);
  o.cmp = take(kDexCmp); o.sums = take(3);
  o.total = k;
  return o;
}
__host__ __device__ inline size_t dex_post_smem_bytes() { return (size_t)dex_stage_ofs().total * 4 + HA_DEX_OBS_DIM * sizeof(HandCol); }

// Descriptor of observation column j (get_num_obs_dict order, :168-199)
__device__ inline HandCol dex_col(const HaDexParams& p, const DexStageOfs& o, int j) {
  auto copy = [](int base, int stride) { return HandCol{base, stride, kOpCopy, 0.f, 0.f, 0.f}; };
  auto scaled = [](int base, int stride, float s) { return HandCol{base, stride, kOpScale, s, 0.f, 0.f}; };
  auto cmp = [&](int at) { return HandCol{o.cmp + at, kDexCmp, kOpCopy, 0.f, 0.f, 0.f}; };
  int k = j;
  if (k < 16) return HandCol{o.dof + 2 * k, 32, kOpUnscale, p.dof_lower[k], p.dof_upper[k], p.dof_range_rcp[k]};
  k -= 16;
  if (k < 16) return cmp(kCmpDofRand + k);
  k -= 16;
  if (k < 16) return copy(o.dof + 2 * k + 1, 32);
  k -= 16;
  if (k < 16) return scaled(o.frc + k, 16, p.force_torque_obs_scale);
  k -= 16;
  if (k < 6) return k < 3 ? copy(o.obj + 7 + k, 13) : scaled(o.obj + 7 + k, 13, p.vel_obs_scale);
  k -= 6;
  if (k < 16) return copy(o.act + k, 16);
  k -= 16;
  if (k < 3) return copy(o.cube + k, 3);
  k -= 3;
  if (k < 1) return copy(o.handp, 1);
  k -= 1;
  if (k < 3) return copy(o.grav + k, 3);
  k -= 3;
  if (k < 52) return copy(o.tip + k, 52);
  k -= 52;
  if (k < 24) return scaled(o.sens + k, 24, p.force_torque_obs_scale);
  k -= 24;
  if (k < 3) return copy(o.frow + k, 3);
  k -= 3;
  if (k < 2) return cmp(kCmpRotDist + k);
  k -= 2;
  if (k < 4) return cmp(kCmpDelay + k);
  k -= 4;
  if (k < 78) {
    if (k < 16) return copy(o.a_as + k, 16);
    if (k < 32) return copy(o.a_aa + k - 16, 16);
    if (k < 39) return copy(o.a_ps + k - 32, 7);
    if (k < 46) return copy(o.a_pa + k - 39, 7);
    if (k < 62) return copy(o.a_ds + k - 46, 16);
    return copy(o.a_da + k - 62, 16);
  }
  k -= 78;
  if (k < 7) return cmp(kCmpObjPose + k);
  k -= 7;
  if (k < 7) return cmp(kCmpGoalPose + k);
  k -= 7;
  if (k < 4) return cmp(kCmpGoalRel + k);
  k -= 4;
  if (k < 7) return cmp(kCmpCam + k);
  k -= 7;
  return cmp(kCmpCamRel + k);
}

__global__ void __launch_bounds__(kDexThreads)
dex_post_kernel(const __grid_constant__ HaDexParams p, const __grid_constant__ HaDexBuffers b, const float* __restrict__ ext_rand, const int mode) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  constexpr int D = kDexD;
  const unsigned FULL = 0xffffffffu;
  const DexStageOfs o = dex_stage_ofs();
  float* stage = reinterpret_cast<float*>(smem_raw);
  HandCol* cols = reinterpret_cast<HandCol*>(stage + o.total);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = p.n_envs, Dp = p.dofs_per_env;
  const int env0 = blockIdx.x * kDexTile;
  const int nvalid = min(kDexTile, n - env0), last = nvalid - 1;
  const bool full = (mode == kDexPostFull);

  DexRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_DEX_POST_RND_STRIDE;
  R.space = 128;
  R.rng.seed = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_SEED];
  R.rng.step = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_STEP];
  const long long frame = b.dynamic[HA_DEX_DYN_FRAME] + 1;                      // self.frame += 1 (:647) precedes the ADR part

  // ---------------- load ----------------
  {
    auto rows = [&](float* dst, const float* src, size_t row_stride, int width) {   // `width` leading floats of one row per env
      for (int j = tid; j < kDexTile * width; j += kDexThreads) {
        const int e = j / width, c = j - width * e;
        cp_async4(dst + j, src + (size_t)min(e, last) * row_stride + c);
      }
    };
    for (int j = tid; j < kDexTile * 13; j += kDexThreads) {                    // object root rows
      const int e = j / 13, c = j - 13 * e, env = env0 + min(e, last);
      const size_t row = p.root_regular ? (size_t)env * p.actors_per_env + p.object_slot : (size_t)__ldg(b.object_actor + env);
      cp_async4(stage + o.obj + j, b.root_state + row * 13 + c);
    }
    const size_t rbs = (size_t)p.n_bodies * 13;
    rows(stage + o.goal, b.goal_states + (size_t)env0 * 13, 13, 7);
    rows(stage + o.palm, b.rigid_body_state + (size_t)env0 * rbs + (size_t)p.palm_body * 13, rbs, 7);
    rows(stage + o.frow, b.rb_forces + ((size_t)env0 * p.n_bodies + p.object_body) * 3, (size_t)p.n_bodies * 3, 3);
    for (int j = tid; j < kDexTile * 52; j += kDexThreads) {                    // fingertip rigid-body rows (all 13 floats)
      const int e = j / 52, r = j - 52 * e, t = r / 13, c = r - 13 * t;
      cp_async4(stage + o.tip + j, b.rigid_body_state + (size_t)(env0 + min(e, last)) * rbs + (size_t)p.tip_body[t] * 13 + c);
    }
    auto tile = [&](float* dst, const float* src, int per_env) {                // contiguous [nvalid, per_env] image
      const float* s0 = src + (size_t)env0 * per_env;
      const int count = nvalid * per_env;
      if (nvalid == kDexTile && ((per_env * kDexTile) & 3) == 0) {
        for (int q = tid; q < (count >> 2); q += kDexThreads) cp_async16(reinterpret_cast<float4*>(dst) + q, reinterpret_cast<const float4*>(s0) + q);
      } else {
        for (int j = tid; j < count; j += kDexThreads) cp_async4(dst + j, s0 + j);
      }
    };
    if (Dp == D) tile(stage + o.dof, b.dof_state, 2 * D);
    else for (int j = tid; j < nvalid * 2 * D; j += kDexThreads) {
      const int e = j / (2 * D), r = j - 2 * D * e;
      cp_async4(stage + o.dof + j, b.dof_state + (size_t)(env0 + e) * 2 * Dp + r);
    }
    tile(stage + o.frc, b.dof_force, 16); tile(stage + o.sens, b.force_sensors, 24); tile(stage + o.act, b.actions, 16);
    tile(stage + o.cur, b.cur_targets, 16); tile(stage + o.prev, b.prev_targets, 16);
    tile(stage + o.a_as, b.affine_actions_scaling, 16); tile(stage + o.a_aa, b.affine_actions_additive, 16);
    tile(stage + o.a_ps, b.affine_cube_pose_scaling, 7); tile(stage + o.a_pa, b.affine_cube_pose_additive, 7);
    tile(stage + o.a_ds, b.affine_dof_pos_scaling, 16); tile(stage + o.a_da, b.affine_dof_pos_additive, 16);
    tile(stage + o.cube, b.cube_random_params, 3); tile(stage + o.handp, b.hand_random_params, 1); tile(stage + o.grav, b.gravity_vec, 3);
    tile(stage + o.freq, b.obs_object_pose_freq, 7); tile(stage + o.cam, b.obs_object_pose, 7);
  }
  cp_async_commit();
  for (int j = tid; j < HA_DEX_OBS_DIM; j += kDexThreads) cols[j] = dex_col(p, o, j);
  // per-env scalars (lane = env) fetched while the copies are in flight
  const bool valid = lane < nvalid;
  const int env = env0 + min(lane, last);
  long long r_reset = 0, r_goal = 0, r_progress = 0, r_hold = 0, r_rate = 1, r_offset = 0, r_lat = 0;
  float r_succ = 0.f, r_best = 0.f, r_pose_white = 0.f, r_obs_delay = 0.f, r_act_delay = 0.f;
  if (warp == 0) {
    r_rate = __ldg(b.cube_pose_refresh_rate + env); r_offset = __ldg(b.cube_pose_refresh_offset + env); r_lat = __ldg(b.action_latency + env);
    r_pose_white = __ldg(b.adr_affine_cube_pose_white + env); r_obs_delay = __ldg(b.adr_cube_obs_delay_prob + env);
    r_act_delay = __ldg(b.adr_action_delay_prob + env);
  } else if (warp == 1) {
    r_best = b.best_rotation_dist[env];
    if (full) {
      r_reset = b.reset_buf[env]; r_goal = b.reset_goal_buf[env]; r_progress = b.progress_buf[env]; r_hold = b.hold_count_buf[env];
      r_succ = b.successes[env];
    }
  }
  cp_async_wait<0>();
  __syncthreads();

  // ---------------- compute ----------------
  float* cmp = stage + o.cmp + lane * kDexCmp;
  float goal_dist = 0.f, rot_dist = 0.f;
  if (warp == 0) {
    const float* ob = stage + o.obj + lane * 13;
    const float* gl = stage + o.goal + lane * 7;
    const float* pl = stage + o.palm + lane * 7;
    const Mat3 Rp = dex_quat_to_matrix(f4{pl[3], pl[4], pl[5], pl[6]});
    const Pose7 op = dex_pose_wrt_wrist(ob, Rp, pl);
    const Pose7 gp = dex_pose_wrt_wrist(gl, Rp, pl);
    const f4 gq_conj = quat_conjugate(f4{gp.v[3], gp.v[4], gp.v[5], gp.v[6]});
    const f4 grel = quat_mul(f4{op.v[3], op.v[4], op.v[5], op.v[6]}, gq_conj);
#pragma unroll
    for (int c = 0; c < 7; ++c) { cmp[kCmpObjPose + c] = op.v[c]; cmp[kCmpGoalPose + c] = gp.v[c]; }
    cmp[kCmpGoalRel + 0] = grel.x; cmp[kCmpGoalRel + 1] = grel.y; cmp[kCmpGoalRel + 2] = grel.z; cmp[kCmpGoalRel + 3] = grel.w;
    // ADR pose pipeline (:1446-1461)
    const bool update_freq = ((frame + r_offset) % r_rate) == 0;
    const float sd = dex_stdev(r_pose_white);
    const float* ps = stage + o.a_ps + lane * 7;
    const float* pa = stage + o.a_pa + lane * 7;
    // the ten uniforms (slots 7..16) and seven normals (slots 0..6) of this env: five counter blocks in fast mode
    const float4 ua = R.u4(env, HA_DEX_POST_RND_CUBE), ub = R.u4(env, HA_DEX_POST_RND_CUBE + 4), uc = R.u4(env, HA_DEX_POST_RND_CUBE + 8);
    const float4 na = R.nrm4(env, HA_DEX_POST_RND_POSE_WHITE), nb = R.nrm4(env, HA_DEX_POST_RND_POSE_WHITE + 4);
    const float pose_white[7] = {na.x, na.y, na.z, na.w, nb.x, nb.y, nb.z};
    const float rf[5] = {2.0f * ua.x + -1.0f, 2.0f * ua.y + -1.0f, 2.0f * ua.z + -1.0f, 2.0f * ua.w + -1.0f, 2.0f * ub.x + -1.0f};
    const f4 rq = dex_new_rot(p.apply_random_quat, rf[3], rf[4], ub.y, ub.z, ub.w);
    const bool inject = uc.x < p.random_cube_pose_prob;
    const bool delayed = uc.y < r_obs_delay;
    const float* init = b.object_init_state + (size_t)env * 13;
    const float rnd_pose[7] = {__ldg(init + 0) + 0.5f * rf[0], __ldg(init + 1) + 0.5f * rf[1], __ldg(init + 2) + 0.5f * rf[2], rq.x, rq.y, rq.z, rq.w};
    float* fr = stage + o.freq + lane * 7;
    float* cm = stage + o.cam + lane * 7;
    float camq[4];
#pragma unroll
    for (int c = 0; c < 7; ++c) {
      const float noisy = ps[c] * op.v[c] + pa[c] + pose_white[c] * sd;
      const float seen = inject ? rnd_pose[c] : noisy;
      const float f = update_freq ? seen : fr[c];
      const float v = delayed ? cm[c] : f;
      fr[c] = f; cm[c] = v;
      cmp[kCmpCam + c] = v;
      if (c >= 3) camq[c - 3] = v;
    }
    const f4 crel = quat_mul(f4{camq[0], camq[1], camq[2], camq[3]}, gq_conj);
    cmp[kCmpCamRel + 0] = crel.x; cmp[kCmpCamRel + 1] = crel.y; cmp[kCmpCamRel + 2] = crel.z; cmp[kCmpCamRel + 3] = crel.w;
    cmp[kCmpDelay + 0] = r_obs_delay; cmp[kCmpDelay + 1] = (float)r_rate / 6.0f; cmp[kCmpDelay + 2] = r_act_delay; cmp[kCmpDelay + 3] = (float)r_lat / 60.0f;
    if (valid) {
#pragma unroll
      for (int c = 0; c < 7; ++c) b.object_pose[(size_t)env * 7 + c] = ob[c];  // the gather the next reset_idx reads (:613, :841)
    }
  } else if (warp == 1) {
    const float* ob = stage + o.obj + lane * 13;
    const float* gl = stage + o.goal + lane * 7;
    const f4 qd = quat_mul(f4{ob[3], ob[4], ob[5], ob[6]}, quat_conjugate(f4{gl[3], gl[4], gl[5], gl[6]}));
    rot_dist = 2.0f * asinf(fminf(norm3(f3{qd.x, qd.y, qd.z}), 1.0f));
    goal_dist = norm3(f3{ob[0] - gl[0], ob[1] - gl[1], ob[2] - gl[2]});
    r_best = r_best < 0.0f ? rot_dist : r_best;                                 // lazy init (:685)
    cmp[kCmpRotDist + 0] = rot_dist; cmp[kCmpRotDist + 1] = r_best;
  } else {
    // warps 2-3: four lanes per env, four DOFs per lane (one Philox block gives the lane's four normals); 8 envs per warp pass
    const int q = lane & 3, sub = lane >> 2, d0 = 4 * q;
    for (int e = (warp - 2) * 8 + sub; e < kDexTile; e += 16) {
      const int ee = env0 + min(e, last);
      const float sd = dex_stdev(__ldg(b.adr_affine_dof_pos_white + ee));
      const float4 wn4 = R.nrm4(ee, HA_DEX_POST_RND_DOF_WHITE + d0);
      const float wn[4] = {wn4.x, wn4.y, wn4.z, wn4.w};
      float s_a = 0.f, s_d = 0.f, s_v = 0.f;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int d = d0 + c, at = e * 16 + d;
        const float lo = p.dof_lower[d], hi = p.dof_upper[d];
        const float dof_pos = stage[o.dof + e * 32 + 2 * d], dof_vel = stage[o.dof + e * 32 + 2 * d + 1];
        const float rz = stage[o.a_ds + at] * dof_pos + stage[o.a_da + at] + wn[c] * sd;
        stage[o.cmp + e * kDexCmp + kCmpDofRand + d] = div_by_const(2.0f * rz - hi - lo, hi - lo, p.dof_range_rcp[d]);
        const float a = stage[o.act + at], dt = stage[o.cur + at] - stage[o.prev + at], vv = dof_vel / (5.0f - 1.0f);
        s_a += a * a; s_d += dt * dt; s_v += vv * vv;
      }
#pragma unroll
      for (int m = 2; m > 0; m >>= 1) {
        s_a += __shfl_xor_sync(FULL, s_a, m); s_d += __shfl_xor_sync(FULL, s_d, m); s_v += __shfl_xor_sync(FULL, s_v, m);
      }
      if (q == 0) { stage[o.sums + e * 3 + 0] = s_a; stage[o.sums + e * 3 + 1] = s_d; stage[o.sums + e * 3 + 2] = s_v; }
    }
  }
  __syncthreads();

  // ---------------- assemble (warps 0, 2, 3) ----------------
  if (warp != 1) {
    const int r = (warp == 0) ? lane : tid - 32;                                // ranks 0..95
    assemble_rows<(HA_DEX_OBS_DIM + 31) / 32>(stage, cols, b.obs_rows + (size_t)env0 * HA_DEX_OBS_DIM, HA_DEX_OBS_DIM, nvalid, p.clip_obs, lane,
                                              warp == 0 ? 0 : warp - 1, 3);
#pragma unroll
    for (int v = 0; v < 2; ++v)                                                  // concatenated policy / critic views (rl_games adapter)
      if (p.n_view_cols[v] > 0)
        assemble_rows<(HA_DEX_OBS_DIM + 31) / 32>(stage, cols, b.view_rows[v] + (size_t)env0 * p.n_view_cols[v], p.n_view_cols[v], nvalid,
                                                  p.clip_obs, lane, warp == 0 ? 0 : warp - 1, 3, b.view_cols[v]);
    for (int j = r; j < nvalid * 7; j += kDexThreads - 32) {                    // the two pose latches
      b.obs_object_pose_freq[(size_t)env0 * 7 + j] = stage[o.freq + j];
      b.obs_object_pose[(size_t)env0 * 7 + j] = stage[o.cam + j];
    }
    if (full)
      for (int j = r; j < nvalid * 16; j += kDexThreads - 32) b.prev_targets[(size_t)env0 * 16 + j] = stage[o.cur + j];   // (:1046)
    return;
  }

  // ---------------- reward + statistics (warp 1, lane = env) ----------------
  if (!full) {
    if (valid) b.best_rotation_dist[env] = r_best;
    if (lane == 0 && atomicAdd(b.tickets + 2, 1u) == gridDim.x - 1) {           // every CTA has read the frame counter by now
      b.dynamic[HA_DEX_DYN_FRAME] = frame;
      b.tickets[2] = 0;
    }
    return;
  }
  long long a_rew = 0, a_fin = 0, a_succ = 0;
  int a_n = 0, a_resets = 0, a_goal = 0, a_timeouts = 0;
  {
    const float* sm = stage + o.sums + lane * 3;
    const long long progress_in = r_progress + 1;                               // progress_buf += 1 (:1035)
    const float dist_rew = goal_dist * p.dist_reward_scale;
    const float rot_rew = 1.0f / (fabsf(rot_dist) + p.rot_eps) * p.rot_reward_scale;
    const float action_penalty = p.action_penalty_scale * sm[0];
    const float action_delta_penalty = p.action_delta_penalty_scale * sm[1];
    const float velocity_penalty = -0.05f * sm[2];
    const bool hit = fabsf(rot_dist) <= p.success_tolerance;
    const long long goal_reached = hit ? 1 : r_goal;
    const long long hold = goal_reached != 0 ? r_hold + 1 : 0;
    const long long goal_resets = hold > (long long)p.num_success_hold_steps ? 1 : r_goal;
    const float succ = r_succ + (float)goal_resets;
    const float reach_goal_rew = (goal_resets == 1) ? p.reach_goal_bonus : 0.0f;
    const bool fell = goal_dist >= p.fall_dist;
    const float fall_rew = fell ? p.fall_penalty : 0.0f;
    long long resets = fell ? 1 : r_reset;
    long long prog = progress_in;
    if (p.max_consecutive_successes > 0) {
      prog = hit ? 0 : prog;
      resets = (succ >= (float)p.max_consecutive_successes) ? 1 : resets;
    }
    const bool timed_out = prog >= (long long)(p.max_episode_length - 1);
    resets = timed_out ? 1 : resets;
    const float timeout_rew = timed_out ? 0.5f * p.fall_penalty : 0.0f;
    const float reward = ((((((dist_rew + rot_rew) + action_penalty) + action_delta_penalty) + velocity_penalty) + reach_goal_rew) + fall_rew) + timeout_rew;
    const bool timeout_flag = timed_out && (resets != 0);                       // adr_vec_task.py:290
    if (valid) {
      b.rew_buf[env] = reward;
      b.reset_buf[env] = resets;
      b.reset_goal_buf[env] = goal_resets;
      b.progress_buf[env] = prog;
      b.hold_count_buf[env] = hold;
      b.successes[env] = succ;
      b.best_rotation_dist[env] = fminf(r_best, rot_dist);                      // (:453)
      b.timeout_buf[env] = timeout_flag ? 1 : 0;
      const float terms[HA_DEX_N_STEP_TERMS] = {dist_rew, rot_rew, action_penalty, action_delta_penalty, velocity_penalty, reach_goal_rew,
                                                fall_rew, timeout_rew};
#pragma unroll
      for (int k = 0; k < HA_DEX_N_STEP_TERMS; ++k) b.episode_cumulative[(size_t)k * n + env] = terms[k];
      a_rew = __float2ll_rn(reward * 65536.0f);
      a_n = 1; a_resets = (int)resets; a_goal = goal_resets != 0; a_timeouts = timeout_flag;
      a_fin = (long long)__float2ll_rn(succ * (float)resets);
      a_succ = (long long)__float2ll_rn(succ);
    }
  }
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
  long long old = 0;
  if (lane < HA_HAND_N_PARTIAL_STATS) old = (long long)atomicAdd(reinterpret_cast<unsigned long long*>(accum + lane), (unsigned long long)val);
  const unsigned dep = __any_sync(FULL, old == 0x7fffffffffffffffLL) ? 1u : 0u;  // ticket depends on every returned atomic
  int is_last = 0;
  if (lane == 0) is_last = (atomicAdd(b.tickets + 0, 1u + dep) == gridDim.x - 1);
  is_last = __shfl_sync(FULL, is_last, 0);
  if (!is_last) return;
  __threadfence();
  float fv = 0.f;
  if (lane < HA_HAND_N_PARTIAL_STATS) {
    const long long raw = (long long)atomicExch(reinterpret_cast<unsigned long long*>(accum + lane), 0ULL);
    fv = (lane == HA_HAND_STAT_SUM_REWARD) ? (float)((double)raw / 65536.0) : (float)raw;
    b.stats[lane] = fv;
  }
  const float cnt = __shfl_sync(FULL, fv, HA_HAND_STAT_N), sum_rew = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_REWARD);
  const float num_resets = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_RESETS), finished = __shfl_sync(FULL, fv, HA_HAND_STAT_SUM_FINISHED_SUCCESSES);
  if (lane == 0) {
    const float cons = b.consecutive_successes[0];                              // (:1657-1660)
    const float av = p.av_factor;
    const float upd = av * finished / num_resets + (1.0f - av) * cons;
    const float out = (num_resets > 0.0f) ? upd : cons;
    b.consecutive_successes[0] = out;
    b.stats[HA_HAND_STAT_MEAN_REWARD] = sum_rew / cnt;
    b.stats[HA_HAND_STAT_CONSECUTIVE_SUCCESSES] = out;
    b.tickets[0] = 0;
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    b.dynamic[HA_DEX_DYN_RNG_STEP] = b.dynamic[HA_DEX_DYN_RNG_STEP] + 1;
    b.dynamic[HA_DEX_DYN_FRAME] = frame;
  }
}

}  // namespace ha
