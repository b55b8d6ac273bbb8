// Random network adversary of the DeXtreme tasks (isaacgymenvs/utils/rna_util.py:37-150, call site
// allegro_hand_dextreme.py:864-879): the hand-written ends of the chain.  The four wide layers in the middle are plain
// fp32 library GEMMs issued by the host side (handarm_b200/rna.py); see include/handarm_b200.h.
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

constexpr int kRnaIn = HA_RNA_IN_DIMS;        // 16 joint angles + cube pose wrt wrist (pos3 quat4)
constexpr int kRnaF1 = HA_RNA_FEATS1;         // 512

// fc1 + ReLU (rna_util.py:126-127) without torch.cat: the 23 inputs of an env are read from dof_state ([N,16,2], pos) and
// object_pose_wrt_wrist ([N,7]).  K = 23 is far too thin for a tensor-core tile: CTA = 32 envs x 512 features, the weight
// matrix (47 KB) and the tile's inputs sit in shared memory, thread = 1 feature column x 32 envs... every weight is read once
// per CTA from smem into registers and reused for the 32 envs.
__global__ void __launch_bounds__(256)
rna_input_layer_kernel(const int n_envs, const float* __restrict__ dof_state, const float* __restrict__ obj_pose, const int pose_stride,
                       const float* __restrict__ w1, const float* __restrict__ b1, float* __restrict__ out,
                       __nv_bfloat16* __restrict__ out_hi, __nv_bfloat16* __restrict__ out_lo) {
  __shared__ __align__(16) float xs[32][24];            // 23 inputs + one zero pad: rows are read as six 128-bit broadcasts
  const int tid = threadIdx.x, env0 = blockIdx.x * 32;
  const int nvalid = min(32, n_envs - env0);
  for (int i = tid; i < 32 * 24; i += 256) {
    const int e = i / 24, c = i - e * 24;
    float v = 0.f;
    if (e < nvalid && c < kRnaIn) v = (c < 16) ? __ldg(dof_state + ((size_t)(env0 + e) * 16 + c) * 2) : __ldg(obj_pose + (size_t)(env0 + e) * pose_stride + (c - 16));
    xs[e][c] = v;
  }
  __syncthreads();
  // thread -> the feature pair (2 tid, 2 tid + 1): both weight rows in registers, loop over the tile's envs; a warp stores 128
  // contiguous bytes per plane and env (bf16x2 per lane)
  static_assert(kRnaF1 == 2 * 256, "one feature pair per thread");
  const int f = 2 * tid;
  float w0[kRnaIn], w1r[kRnaIn];
#pragma unroll
  for (int c = 0; c < kRnaIn; ++c) { w0[c] = __ldg(w1 + (size_t)f * kRnaIn + c); w1r[c] = __ldg(w1 + (size_t)(f + 1) * kRnaIn + c); }
  const float bias0 = __ldg(b1 + f), bias1 = __ldg(b1 + f + 1);
  for (int e = 0; e < nvalid; ++e) {
    float a0 = 0.f, a1 = 0.f;
    float x[24];
#pragma unroll
    for (int c4 = 0; c4 < 6; ++c4) {
      const float4 t = *reinterpret_cast<const float4*>(&xs[e][4 * c4]);
      x[4 * c4] = t.x; x[4 * c4 + 1] = t.y; x[4 * c4 + 2] = t.z; x[4 * c4 + 3] = t.w;
    }
#pragma unroll
    for (int c = 0; c < kRnaIn; ++c) { a0 = fmaf(x[c], w0[c], a0); a1 = fmaf(x[c], w1r[c], a1); }
    const float v0 = fmaxf(a0 + bias0, 0.f), v1 = fmaxf(a1 + bias1, 0.f);
    const size_t o = (size_t)(env0 + e) * kRnaF1 + f;
    if (out_hi) {                                                         // operand planes of the tensor-core layers
      const __nv_bfloat16 h0 = __float2bfloat16_rn(v0), h1 = __float2bfloat16_rn(v1);
      __nv_bfloat162 hh, ll;
      hh.x = h0; hh.y = h1;
      ll.x = __float2bfloat16_rn(v0 - __bfloat162float(h0)); ll.y = __float2bfloat16_rn(v1 - __bfloat162float(h1));
      *reinterpret_cast<__nv_bfloat162*>(out_hi + o) = hh;
      *reinterpret_cast<__nv_bfloat162*>(out_lo + o) = ll;
    } else {
      *reinterpret_cast<float2*>(out + o) = make_float2(v0, v1);
    }
  }
}

// Operand planes of fc1 for the tensor-core path: row = [dof_pos 16 | object pose 7 | zeros to HA_RNA_IN_PAD], split to bf16 hi / lo.
// One thread per (env, column); 4 MB written per step at 16384 envs.
__global__ void __launch_bounds__(256)
rna_input_planes_kernel(const int n_envs, const float* __restrict__ dof_state, const float* __restrict__ obj_pose, const int pose_stride,
                        __nv_bfloat16* __restrict__ out_hi, __nv_bfloat16* __restrict__ out_lo) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)n_envs * HA_RNA_IN_PAD) return;
  const int env = (int)(i / HA_RNA_IN_PAD), c = (int)(i - (size_t)env * HA_RNA_IN_PAD);
  float v = 0.f;
  if (c < 16) v = __ldg(dof_state + ((size_t)env * 16 + c) * 2);
  else if (c < kRnaIn) v = __ldg(obj_pose + (size_t)env * pose_stride + (c - 16));
  const __nv_bfloat16 h = __float2bfloat16_rn(v);
  out_hi[i] = h;
  out_lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
}

// x = (x + bias) [relu] [* mask], in place, 128-bit accesses; cols % 4 == 0.
// fc1_1 / fc2_1 + dropout mask (rna_util.py:128-129, :133-134: no activation before the mask), fc2 + ReLU (:131-132).
__global__ void __launch_bounds__(256)
rna_bias_act_kernel(float* __restrict__ x, const float* __restrict__ bias, const float* __restrict__ mask, const int relu,
                    const size_t nvec, const int cols4) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < nvec; q += stride) {
    const int c = (int)(q % (size_t)cols4);
    float4 v = reinterpret_cast<float4*>(x)[q];
    const float4 bv = __ldg(reinterpret_cast<const float4*>(bias) + c);
    v.x += bv.x; v.y += bv.y; v.z += bv.z; v.w += bv.w;
    if (relu) { v.x = fmaxf(v.x, 0.f); v.y = fmaxf(v.y, 0.f); v.z = fmaxf(v.z, 0.f); v.w = fmaxf(v.w, 0.f); }
    if (mask) {
      const float4 m = __ldg(reinterpret_cast<const float4*>(mask) + q);
      v.x = m.x * v.x; v.y = m.y * v.y; v.z = m.z * v.z; v.w = m.w * v.w;
    }
    reinterpret_cast<float4*>(x)[q] = v;
  }
}

// fc3 bias + softmax + argmax + gather (rna_util.py:136-139, allegro_hand_dextreme.py:871-876): softmax is monotone, so the bin
// is the argmax of the logits (first maximum, as torch.argmax); out[env, d] = discretised_dofs[d, bin].  Warp = env, lane = bin.
__global__ void __launch_bounds__(256)
rna_select_kernel(const int n_envs, const float* __restrict__ logits, const float* __restrict__ bias3,
                  const float* __restrict__ discretised, float* __restrict__ out, int32_t* __restrict__ idx_out) {
  const int lane = threadIdx.x & 31;
  const int env = blockIdx.x * (256 / 32) + (threadIdx.x >> 5);
  if (env >= n_envs) return;
  const float* row = logits + (size_t)env * (HA_RNA_DOFS * HA_RNA_BINS);
  float res = 0.f;
  int res_idx = 0;
#pragma unroll
  for (int d = 0; d < HA_RNA_DOFS; ++d) {
    float v = row[d * HA_RNA_BINS + lane] + __ldg(bias3 + d * HA_RNA_BINS + lane);
    int idx = lane;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, v, o);
      const int oi = __shfl_xor_sync(0xffffffffu, idx, o);
      if (ov > v || (ov == v && oi < idx)) { v = ov; idx = oi; }
    }
    if (lane == d) { res = __ldg(discretised + d * HA_RNA_BINS + idx); res_idx = idx; }
  }
  if (lane < HA_RNA_DOFS) {
    out[(size_t)env * HA_RNA_DOFS + lane] = res;
    if (idx_out) idx_out[(size_t)env * HA_RNA_DOFS + lane] = res_idx;
  }
}

}  // namespace ha

extern "C" {

int ha_rna_input_layer(int32_t n_envs, const float* dof_state, const float* object_pose_wrt_wrist, int32_t pose_stride, const float* w1, const float* b1,
                       float* out, void* out_hi, void* out_lo, ha_stream_t stream) {
  HA_REQUIRE(n_envs > 0, "n_envs must be positive (got %d)", n_envs);
  HA_REQUIRE(dof_state && object_pose_wrt_wrist && w1 && b1 && (out || (out_hi && out_lo)), "null pointer");
  HA_REQUIRE(pose_stride >= 7, "pose_stride must be >= 7 floats (got %d)", pose_stride);
  const int grid = (n_envs + 31) / 32;
  return ha::check_cuda(ha::launch(ha::rna_input_layer_kernel, grid, 256, 0, (cudaStream_t)stream, (int)n_envs, dof_state, object_pose_wrt_wrist,
                                   (int)pose_stride, w1, b1, out, reinterpret_cast<__nv_bfloat16*>(out_hi), reinterpret_cast<__nv_bfloat16*>(out_lo)),
                        "rna_input_layer_kernel launch");
}

int ha_rna_input_planes(int32_t n_envs, const float* dof_state, const float* object_pose_wrt_wrist, int32_t pose_stride, void* out_hi, void* out_lo,
                        ha_stream_t stream) {
  HA_REQUIRE(n_envs > 0, "n_envs must be positive (got %d)", n_envs);
  HA_REQUIRE(dof_state && object_pose_wrt_wrist && out_hi && out_lo, "null pointer");
  HA_REQUIRE(pose_stride >= 7, "pose_stride must be >= 7 floats (got %d)", pose_stride);
  const size_t count = (size_t)n_envs * HA_RNA_IN_PAD;
  return ha::check_cuda(ha::launch(ha::rna_input_planes_kernel, (int)((count + 255) / 256), 256, 0, (cudaStream_t)stream, (int)n_envs, dof_state,
                                   object_pose_wrt_wrist, (int)pose_stride, reinterpret_cast<__nv_bfloat16*>(out_hi), reinterpret_cast<__nv_bfloat16*>(out_lo)),
                        "rna_input_planes_kernel launch");
}

int ha_rna_bias_act(float* x, const float* bias, const float* mask, int32_t relu, size_t rows, int32_t cols, ha_stream_t stream) {
  HA_REQUIRE(x && bias, "null pointer");
  HA_REQUIRE(cols > 0 && cols % 4 == 0, "cols must be a positive multiple of 4 (got %d)", cols);
  HA_REQUIRE_ALIGNED16(x);
  HA_REQUIRE_ALIGNED16(bias);
  if (mask) HA_REQUIRE_ALIGNED16(mask);
  if (rows == 0) return HA_OK;
  const size_t nvec = rows * (size_t)(cols / 4);
  const int sm_count = ha::device_sm_count();
  const size_t want = (nvec + 255) / 256;
  const int grid = (int)(want < (size_t)sm_count * 8 ? want : (size_t)sm_count * 8);
  return ha::check_cuda(ha::launch(ha::rna_bias_act_kernel, grid, 256, 0, (cudaStream_t)stream, x, bias, mask, (int)relu, nvec, (int)(cols / 4)),
                        "rna_bias_act_kernel launch");
}

int ha_rna_select(int32_t n_envs, const float* logits, const float* bias3, const float* discretised_dofs, float* out, int32_t* idx_out,
                  ha_stream_t stream) {
  HA_REQUIRE(n_envs > 0, "n_envs must be positive (got %d)", n_envs);
  HA_REQUIRE(logits && bias3 && discretised_dofs && out, "null pointer");
  const int grid = (n_envs + 7) / 8;
  return ha::check_cuda(ha::launch(ha::rna_select_kernel, grid, 256, 0, (cudaStream_t)stream, (int)n_envs, logits, bias3, discretised_dofs, out,
                                   idx_out), "rna_select_kernel launch");
}

}  // extern "C"
