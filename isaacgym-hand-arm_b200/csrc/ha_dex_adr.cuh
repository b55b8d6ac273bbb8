// Per-env ADR resampling of AllegroHandDextremeADR on reset: recycle_envs (adr_vec_task.py:740-758), sample_adr_tensor for every
// tensorised parameter (:666-730) and the Dextreme draws of allegro_hand_dextreme.py:1392-1416 (action latency, cube-pose refresh
// rate / offset, the six affine tensors), for the envs in reset_buf & ~apply_reset_buf -- the work is sparse:
// every warp scans 32 envs and then serves its masked envs one at a time, cooperatively (lane = ADR parameter, lane = affine column).
#pragma once
#include "ha_dex_common.cuh"

namespace ha {

__global__ void __launch_bounds__(kDexThreads)
dex_adr_kernel(const __grid_constant__ HaDexParams p, const __grid_constant__ HaDexBuffers b, const __grid_constant__ HaDexAdrTable t,
               const __grid_constant__ HaDexAdrBuffers ab, const float* __restrict__ ext_rand) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const int env_base = (blockIdx.x * kDexThreads + threadIdx.x) & ~31;
  if (env_base >= p.n_envs) return;
  const int my_env = env_base + lane;
  const bool mine = my_env < p.n_envs && (t.all_envs || (b.reset_buf[my_env] & ~b.apply_reset_buf[my_env]) != 0);      // env_mask_randomize (:930)
  unsigned todo = __ballot_sync(FULL, mine);
  if (todo == 0u) return;
  DexRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_DEX_ADR_RND_STRIDE;
  R.space = 256;
  R.rng.seed = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_SEED];
  R.rng.step = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_STEP];
  float* const dst[6] = {b.affine_actions_scaling, b.affine_actions_additive, b.affine_cube_pose_scaling, b.affine_cube_pose_additive,
                         b.affine_dof_pos_scaling, b.affine_dof_pos_additive};
  const int n_modes = 2 * t.n_params_total;
  while (todo) {
    const int env = env_base + __ffs(todo) - 1;
    todo &= todo - 1;
    // recycle_envs (every lane derives the same worker type and mode)
    const long long worker = R.u(env, HA_DEX_ADR_RND_WORKER) < t.boundary_fraction ? 0 : 1;      // ADR_ROLLOUT : ADR_BOUNDARY
    const long long mode = min((long long)(R.u(env, HA_DEX_ADR_RND_MODE) * (float)n_modes), (long long)n_modes - 1);
    // sample_adr_tensor, lane = parameter: boundary workers evaluate one end of one parameter, everybody else draws from the range
    float v = 0.f;
    if (lane < t.n_tensor) {
      const int n = t.param_index[lane];
      if (worker == 1 && mode == 2 * n) v = t.low_value[lane];
      else if (worker == 1 && mode == 2 * n + 1) v = t.high_value[lane];
      else v = R.u(env, HA_DEX_ADR_RND_PARAM + lane) * t.span[lane] + t.lo[lane];
      ab.values[lane][env] = v;
    }
    const float latency_v = __shfl_sync(FULL, v, t.idx_action_latency), rate_v = __shfl_sync(FULL, v, t.idx_cube_pose_refresh_rate);
    float sd[6];
#pragma unroll
    for (int a = 0; a < 6; ++a) sd[a] = dex_stdev(__shfl_sync(FULL, v, t.idx_affine[a]));
    if (lane == 0) {
      ab.worker_types[env] = worker;
      ab.adr_modes[env] = mode;
      // sample_discrete_adr (:1373-1381): round(value + (0.5 - u)), half to even like torch.round
      const long long latency = (long long)rintf(latency_v + (-(R.u(env, HA_DEX_ADR_RND_LATENCY) - 0.5f)));
      const long long rate = (long long)rintf(rate_v + (-(R.u(env, HA_DEX_ADR_RND_RATE) - 0.5f)));
      b.action_latency[env] = latency;
      b.cube_pose_refresh_rate[env] = rate;
      b.cube_pose_refresh_offset[env] = (long long)rintf(R.u(env, HA_DEX_ADR_RND_OFFSET) * (float)rate - 0.5f);      // (:1404-1406)
    }
    // the six affine tensors (:1408-1416), lane = column of the 78-wide concatenation: [1 +] N(0,1) * stdev(adr value)
    for (int j = lane; j < 78; j += 32) {
      const int a = j < 16 ? 0 : j < 32 ? 1 : j < 39 ? 2 : j < 46 ? 3 : j < 62 ? 4 : 5;
      const int start = a == 0 ? 0 : a == 1 ? 16 : a == 2 ? 32 : a == 3 ? 39 : a == 4 ? 46 : 62;
      const int width = (a == 2 || a == 3) ? 7 : kDexD;
      float s_a = sd[0];
#pragma unroll
      for (int k = 1; k < 6; ++k) s_a = (a == k) ? sd[k] : s_a;
      const float noise = R.nrm(env, HA_DEX_ADR_RND_AFFINE + j) * s_a;
      float* out = a == 0 ? dst[0] : a == 1 ? dst[1] : a == 2 ? dst[2] : a == 3 ? dst[3] : a == 4 ? dst[4] : dst[5];
      out[(size_t)env * width + (j - start)] = (a & 1) == 0 ? 1.0f + noise : noise;
    }
  }
}

}  // namespace ha
