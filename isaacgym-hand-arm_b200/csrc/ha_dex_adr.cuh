// Per-env ADR resampling of AllegroHandDextremeADR on reset: recycle_envs (adr_vec_task.py:740-758), sample_adr_tensor for every
// tensorised parameter (:666-730) and the Dextreme draws of allegro_hand_dextreme.py:1392-1416 (action latency, cube-pose refresh
// rate / offset, the six affine tensors), for the envs in reset_buf & ~apply_reset_buf -- one thread per env, the work is sparse.
#pragma once
#include "ha_dex_common.cuh"

namespace ha {

__global__ void __launch_bounds__(kDexThreads)
dex_adr_kernel(const __grid_constant__ HaDexParams p, const __grid_constant__ HaDexBuffers b, const __grid_constant__ HaDexAdrTable t,
               const __grid_constant__ HaDexAdrBuffers ab, const float* __restrict__ ext_rand) {
  const int env = blockIdx.x * kDexThreads + threadIdx.x;
  if (env >= p.n_envs) return;
  if (!t.all_envs && (b.reset_buf[env] & ~b.apply_reset_buf[env]) == 0) return;       // env_mask_randomize (:930)
  DexRand R;
  R.ext = p.use_fast_rng ? nullptr : ext_rand;
  R.stride = HA_DEX_ADR_RND_STRIDE;
  R.space = 256;
  R.rng.seed = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_SEED];
  R.rng.step = (uint64_t)b.dynamic[HA_DEX_DYN_RNG_STEP];
  // recycle_envs
  const long long worker = R.u(env, HA_DEX_ADR_RND_WORKER) < t.boundary_fraction ? 0 : 1;      // ADR_ROLLOUT : ADR_BOUNDARY
  const int n_modes = 2 * t.n_params_total;
  const long long mode = min((long long)(R.u(env, HA_DEX_ADR_RND_MODE) * (float)n_modes), (long long)n_modes - 1);
  ab.worker_types[env] = worker;
  ab.adr_modes[env] = mode;
  // sample_adr_tensor: boundary workers evaluate one end of one parameter, everybody else draws from the current range
  float latency_v = 0.f, rate_v = 0.f, affine_v[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int k = 0; k < t.n_tensor; ++k) {
    const int n = t.param_index[k];
    float v;
    if (worker == 1 && mode == 2 * n) v = t.low_value[k];
    else if (worker == 1 && mode == 2 * n + 1) v = t.high_value[k];
    else v = R.u(env, HA_DEX_ADR_RND_PARAM + k) * t.span[k] + t.lo[k];
    ab.values[k][env] = v;
    if (k == t.idx_action_latency) latency_v = v;
    if (k == t.idx_cube_pose_refresh_rate) rate_v = v;
#pragma unroll
    for (int a = 0; a < 6; ++a)
      if (k == t.idx_affine[a]) affine_v[a] = v;
  }
  // sample_discrete_adr (:1373-1381): round(value + (0.5 - u)), half to even like torch.round
  const long long latency = (long long)rintf(latency_v + (-(R.u(env, HA_DEX_ADR_RND_LATENCY) - 0.5f)));
  const long long rate = (long long)rintf(rate_v + (-(R.u(env, HA_DEX_ADR_RND_RATE) - 0.5f)));
  b.action_latency[env] = latency;
  b.cube_pose_refresh_rate[env] = rate;
  b.cube_pose_refresh_offset[env] = (long long)rintf(R.u(env, HA_DEX_ADR_RND_OFFSET) * (float)rate - 0.5f);      // (:1404-1406)
  // the six affine tensors (:1408-1416): [1 +] N(0,1) * stdev(adr value)
  float* dst[6] = {b.affine_actions_scaling, b.affine_actions_additive, b.affine_cube_pose_scaling, b.affine_cube_pose_additive,
                   b.affine_dof_pos_scaling, b.affine_dof_pos_additive};
  const int dims[6] = {kDexD, kDexD, 7, 7, kDexD, kDexD};
  int slot = HA_DEX_ADR_RND_AFFINE;
#pragma unroll
  for (int a = 0; a < 6; ++a) {
    const float sd = dex_stdev(affine_v[a]);
    const bool scaling = (a & 1) == 0;
    for (int j = 0; j < dims[a]; ++j) {
      const float noise = R.nrm(env, slot + j) * sd;
      dst[a][(size_t)env * dims[a] + j] = scaling ? 1.0f + noise : noise;
    }
    slot += dims[a];
  }
}

}  // namespace ha
