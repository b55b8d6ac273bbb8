// Shared device code of the DeXtreme (AllegroHandDextremeADR) kernels: random-number access (external table in parity
// mode, Philox otherwise), sample_gaussian_adr's standard deviation, and the wrist-relative pose of
// compute_poses_wrt_wrist (allegro_hand_dextreme.py:533-588) with the wxyz matrix conversions of torch_jit_utils.py:462-544.
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"

namespace ha {

constexpr int kDexThreads = 128;
constexpr int kDexTile = 32;
constexpr int kDexD = HA_DEX_DOFS;

struct DexRand {
  const float* ext;   // [n_envs, stride] or nullptr (fast mode)
  int stride;
  int space;          // separates the Philox counters of the pre and the post kernel
  Rng rng;
  __device__ __forceinline__ float u(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.uniform(env, space + slot);
  }
  __device__ __forceinline__ float nrm(int env, int slot) const {
    return ext ? __ldg(ext + (size_t)env * stride + slot) : rng.normal(env, space + slot);
  }
  // normals of slots slot .. slot+3: the table's four entries, or one Philox block in fast mode
  __device__ __forceinline__ float4 nrm4(int env, int slot) const {
    if (ext) {
      const float* r = ext + (size_t)env * stride + slot;
      return make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
    }
    return rng.normal4(env, (uint32_t)(space + slot));
  }
  // the three normals slot .. slot+2 of one env: the table's entries, or ONE counter block in fast mode
  __device__ __forceinline__ void nrm3(int env, int slot, float (&o)[3]) const {
    if (ext) {
      const float* r = ext + (size_t)env * stride + slot;
      o[0] = __ldg(r); o[1] = __ldg(r + 1); o[2] = __ldg(r + 2);
    } else {
      rng.normal3(env, (uint32_t)(space + slot), o);
    }
  }
  // uniforms of slots slot .. slot+3
  __device__ __forceinline__ float4 u4(int env, int slot) const {
    if (ext) {
      const float* r = ext + (size_t)env * stride + slot;
      return make_float4(__ldg(r), __ldg(r + 1), __ldg(r + 2), __ldg(r + 3));
    }
    return rng.uniform4(env, 0x20000000 | (space + slot));
  }
};

// sample_gaussian_adr (:1383-1387): stdev = exp(v^2) - 1 where v > 0, else 0
__device__ __forceinline__ float dex_stdev(float v) { return v > 0.0f ? expf(v * v) - 1.0f : 0.0f; }

struct Mat3 { float m[3][3]; };

// quaternion_to_matrix (torch_jit_utils.py:462-486) of an xyzw quaternion (the reference reorders to wxyz first)
__device__ __forceinline__ Mat3 dex_quat_to_matrix(const f4 q) {
  const float r = q.w, i = q.x, j = q.y, k = q.z;
  const float two_s = 2.0f / (((r * r + i * i) + j * j) + k * k);
  Mat3 R;
  R.m[0][0] = 1.0f - two_s * (j * j + k * k); R.m[0][1] = two_s * (i * j - k * r);        R.m[0][2] = two_s * (i * k + j * r);
  R.m[1][0] = two_s * (i * j + k * r);        R.m[1][1] = 1.0f - two_s * (i * i + k * k); R.m[1][2] = two_s * (j * k - i * r);
  R.m[2][0] = two_s * (i * k - j * r);        R.m[2][1] = two_s * (j * k + i * r);        R.m[2][2] = 1.0f - two_s * (i * i + j * j);
  return R;
}

// matrix_to_quaternion (torch_jit_utils.py:501-544): best-conditioned candidate, returned xyzw
__device__ __forceinline__ f4 dex_matrix_to_quat(const Mat3& M) {
  const float m00 = M.m[0][0], m01 = M.m[0][1], m02 = M.m[0][2], m10 = M.m[1][0], m11 = M.m[1][1], m12 = M.m[1][2];
  const float m20 = M.m[2][0], m21 = M.m[2][1], m22 = M.m[2][2];
  float x[4] = {1.0f + m00 + m11 + m22, 1.0f + m00 - m11 - m22, 1.0f - m00 + m11 - m22, 1.0f - m00 - m11 + m22};
  float qa[4];
  int best = 0;
#pragma unroll
  for (int c = 0; c < 4; ++c) {
    qa[c] = x[c] > 0.0f ? sqrtf(x[c]) : 0.0f;
    if (qa[c] > qa[best]) best = c;              // first maximum, as argmax
  }
  float r, i, j, k;                               // candidate row `best` of quat_by_rijk (wxyz)
  if (best == 0)      { r = qa[0] * qa[0]; i = m21 - m12;       j = m02 - m20;       k = m10 - m01; }
  else if (best == 1) { r = m21 - m12;     i = qa[1] * qa[1];   j = m10 + m01;       k = m02 + m20; }
  else if (best == 2) { r = m02 - m20;     i = m10 + m01;       j = qa[2] * qa[2];   k = m12 + m21; }
  else                { r = m10 - m01;     i = m20 + m02;       j = m21 + m12;       k = qa[3] * qa[3]; }
  const float den = 2.0f * fmaxf(qa[best], 0.1f);
  return f4{i / den, j / den, k / den, r / den};
}

struct Pose7 { float v[7]; };

// relative pose inverse(T_palm) @ T_pose: rigid inverse [R^T | -R^T p] followed by the 4x4 product in k order
__device__ __forceinline__ Pose7 dex_pose_wrt_wrist(const float* pose, const Mat3& Rp, const float* palm_pos) {
  float ti[3];                                    // last column of the inverse: -(R^T p)
#pragma unroll
  for (int r = 0; r < 3; ++r) ti[r] = -((Rp.m[0][r] * palm_pos[0] + Rp.m[1][r] * palm_pos[1]) + Rp.m[2][r] * palm_pos[2]);
  const Mat3 Ro = dex_quat_to_matrix(f4{pose[3], pose[4], pose[5], pose[6]});
  Mat3 Rr;
  Pose7 out;
#pragma unroll
  for (int r = 0; r < 3; ++r) {
#pragma unroll
    for (int c = 0; c < 3; ++c) Rr.m[r][c] = (Rp.m[0][r] * Ro.m[0][c] + Rp.m[1][r] * Ro.m[1][c]) + Rp.m[2][r] * Ro.m[2][c];
    out.v[r] = ((Rp.m[0][r] * pose[0] + Rp.m[1][r] * pose[1]) + Rp.m[2][r] * pose[2]) + ti[r];
  }
  const f4 q = dex_matrix_to_quat(Rr);
  out.v[3] = q.x; out.v[4] = q.y; out.v[5] = q.z; out.v[6] = q.w;
  return out;
}

// quat_from_angle_axis about a unit axis, randomize_rotation (:1666-1669) -- the biased sampler used when apply_random_quat is off
__device__ __forceinline__ f4 dex_quat_axis_angle(float angle, int axis) {
  const float th = angle / 2.0f;
  const float s = sinf(th), c = cosf(th);
  const float nrm = fmaxf(sqrtf(s * s + c * c), 1e-9f);
  f4 q = {0.f, 0.f, 0.f, c / nrm};
  const float v = s / nrm;
  if (axis == 0) q.x = v; else q.y = v;
  return q;
}
__device__ __forceinline__ f4 dex_new_rot(const bool random_quat_mode, float r0, float r1, float u, float v, float w) {
  if (random_quat_mode) return random_quat(u, v, w);                        // get_random_quat (:695-707)
  const float pi = 3.1415927410125732f;
  return quat_mul(dex_quat_axis_angle(r0 * pi, 0), dex_quat_axis_angle(r1 * pi, 1));
}

}  // namespace ha
