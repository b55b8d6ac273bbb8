// Observation / action noise of domain randomisation (vec_task.py:684-754) as one streaming kernel; see include/handarm_b200.h.
#include <cuda_runtime.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

// thread = 4 consecutive elements: 128-bit loads / stores, one Philox block for the four white draws
__global__ void __launch_bounds__(256)
dr_noise_kernel(const __grid_constant__ HaNoiseParams p, const float* x, const float* __restrict__ corr, float* out,      // x may alias out (in place)
                const size_t count, const float* __restrict__ ext_white) {
  Rng rng;
  rng.seed = p.rng_seed;
  rng.step = p.rng_step;
  const bool gaussian = p.distribution == HA_NOISE_GAUSSIAN, additive = p.operation == HA_NOISE_ADDITIVE, clip = p.clip > 0.0f;
  auto one = [&](float xv, float cv, float wv) {
    const float c = cv * p.corr_scale + p.corr_shift;
    const float noise = c + wv * p.white_scale + p.white_shift;
    float v = additive ? xv + noise : xv * noise;
    if (clip) v = fminf(fmaxf(v, -p.clip), p.clip);
    return v;
  };
  const size_t nvec = count >> 2, stride = (size_t)gridDim.x * blockDim.x;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
    const float4 xv = reinterpret_cast<const float4*>(x)[q], cv = __ldg(reinterpret_cast<const float4*>(corr) + q);
    float4 w;
    if (ext_white) w = __ldg(reinterpret_cast<const float4*>(ext_white) + q);
    else {
      const int lo = (int)(q & 0x7fffffffu);
      const uint32_t hi = (uint32_t)(q >> 31);
      w = gaussian ? rng.normal4(lo, hi) : rng.uniform4(lo, (int)(0x10000000u | hi));
    }
    reinterpret_cast<float4*>(out)[q] = make_float4(one(xv.x, cv.x, w.x), one(xv.y, cv.y, w.y), one(xv.z, cv.z, w.z), one(xv.w, cv.w, w.w));
  }
  const size_t tail = (nvec << 2) + (size_t)blockIdx.x * blockDim.x + threadIdx.x;      // up to 3 trailing elements
  if (tail < count) {
    float w;
    if (ext_white) w = __ldg(ext_white + tail);
    else w = gaussian ? rng.normal((int)(tail & 0x7fffffffu), 7) : rng.uniform((int)(tail & 0x7fffffffu), 7);
    out[tail] = one(x[tail], __ldg(corr + tail), w);
  }
}

}  // namespace ha

extern "C" {

size_t ha_sizeof_noise_params(void) { return sizeof(HaNoiseParams); }

int ha_dr_noise_apply(const HaNoiseParams* p, const float* x, const float* corr, float* out, size_t count, const float* ext_white,
                      ha_stream_t stream) {
  HA_REQUIRE(p != nullptr, "null params");
  HA_REQUIRE(p->abi_version == HA_ABI_VERSION, "HaNoiseParams.abi_version %d != %d", p->abi_version, HA_ABI_VERSION);
  HA_REQUIRE(p->distribution == HA_NOISE_GAUSSIAN || p->distribution == HA_NOISE_UNIFORM, "unknown distribution %d", p->distribution);
  HA_REQUIRE(p->operation == HA_NOISE_ADDITIVE || p->operation == HA_NOISE_SCALING, "unknown operation %d", p->operation);
  HA_REQUIRE(x != nullptr && corr != nullptr && out != nullptr, "x / corr / out is null");
  HA_REQUIRE(p->use_fast_rng || ext_white != nullptr, "ext_white is required unless use_fast_rng is set");
  HA_REQUIRE_ALIGNED16(x);
  HA_REQUIRE_ALIGNED16(corr);
  HA_REQUIRE_ALIGNED16(out);
  if (ext_white) HA_REQUIRE_ALIGNED16(ext_white);
  if (count == 0) return HA_OK;
  const int sm_count = ha::device_sm_count();
  const size_t want = ((count >> 2) + 255) / 256 + 1;
  const int grid = (int)(want < (size_t)sm_count * 8 ? want : (size_t)sm_count * 8);      // persistent: 8 CTAs of 256 threads per SM
  return ha::check_cuda(ha::launch(ha::dr_noise_kernel, grid, 256, 0, (cudaStream_t)stream, *p, x, corr, out, count,
                                   p->use_fast_rng ? (const float*)nullptr : ext_white), "dr_noise_kernel launch");
}

}  // extern "C"
