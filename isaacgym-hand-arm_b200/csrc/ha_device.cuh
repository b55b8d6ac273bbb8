// Device-side building blocks shared by the task kernels (sm_100a).
//
// Arithmetic contract: this translation unit is compiled with -fmad=false and
// every helper spells out the association order of the reference's eager
// PyTorch ops (isaacgymenvs/utils/torch_jit_utils.py), so +,-,*,/ and sqrt are
// bit-identical to the CPU reference; only transcendental functions (log, exp,
// sin, cos) may differ in the last ulp.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ha {

constexpr int kEnvsPerCta = 32;   // == HA_ENVS_PER_CTA
constexpr int kThreads = 128;     // 4 warps: lane = env in tile, warp = work part

struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };   // quaternion xyzw

__device__ __forceinline__ f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
__device__ __forceinline__ f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }

// torch_jit_utils.py:80-90 quat_rotate:  a = v*(2w^2-1);  b = cross(q,v)*w*2;
// c = q*(q.v)*2;  a+b+c
__device__ __forceinline__ f3 quat_rotate(f4 q, f3 v) {
  const float s = 2.0f * (q.w * q.w) - 1.0f;
  const f3 a = {v.x * s, v.y * s, v.z * s};
  const f3 cr = {q.y * v.z - q.z * v.y, q.z * v.x - q.x * v.z, q.x * v.y - q.y * v.x};
  const f3 b = {cr.x * q.w * 2.0f, cr.y * q.w * 2.0f, cr.z * q.w * 2.0f};
  const float d = q.x * v.x + q.y * v.y + q.z * v.z;
  const f3 c = {q.x * d * 2.0f, q.y * d * 2.0f, q.z * d * 2.0f};
  return {a.x + b.x + c.x, a.y + b.y + c.y, a.z + b.z + c.z};
}

// torch_jit_utils.py:41-62 quat_mul (8-multiply factored form)
__device__ __forceinline__ f4 quat_mul(f4 a, f4 b) {
  const float ww = (a.z + a.x) * (b.x + b.y);
  const float yy = (a.w - a.y) * (b.w + b.z);
  const float zz = (a.w + a.y) * (b.w - b.z);
  const float xx = ww + yy + zz;
  const float qq = 0.5f * (xx + (a.z - a.x) * (b.x - b.y));
  f4 r;
  r.w = qq - ww + (a.z - a.y) * (b.y - b.z);
  r.x = qq - xx + (a.x + a.w) * (b.x + b.w);
  r.y = qq - yy + (a.w - a.x) * (b.y + b.z);
  r.z = qq - zz + (a.z + a.y) * (b.w - b.x);
  return r;
}

// torch_jit_utils.py:106-110
__device__ __forceinline__ f4 quat_conjugate(f4 a) { return {-a.x, -a.y, -a.z, a.w}; }

__device__ __forceinline__ float norm3(f3 v) { return sqrtf(v.x * v.x + v.y * v.y + v.z * v.z); }

// torch_jit_utils.py:233-240
__device__ __forceinline__ float scale_to(float x, float lo, float hi) { return 0.5f * (x + 1.0f) * (hi - lo) + lo; }
__device__ __forceinline__ float unscale_from(float x, float lo, float hi) { return (2.0f * x - hi - lo) / (hi - lo); }
// torch_jit_utils.py:228-230 tensor_clamp = max(min(t, hi), lo)
__device__ __forceinline__ float clamp_to(float t, float lo, float hi) { return fmaxf(fminf(t, hi), lo); }
// torch.clip(x, lo, hi) = min(max(x, lo), hi)
__device__ __forceinline__ float clipf(float x, float lo, float hi) { return fminf(fmaxf(x, lo), hi); }

// get_random_quat (allegro_kuka_base.py:1178-1189): uniform SO(3) from 3 uniforms
__device__ __forceinline__ f4 random_quat(float u, float v, float w) {
  const float two_pi = 6.2831854820251465f;  // float32(2*pi)
  const float s1 = sqrtf(1.0f - u), s2 = sqrtf(u);
  float sv, cv, sw, cw;
  sincosf(two_pi * v, &sv, &cv);          // one range reduction per angle; same values as sinf / cosf
  sincosf(two_pi * w, &sw, &cw);
  f4 q;
  q.w = s1 * sv;
  q.x = s1 * cv;
  q.y = s2 * sw;
  q.z = s2 * cw;
  return q;
}

// ---------------------------------------------------------------------------
// Philox4x32-10 counter-based RNG (fast mode): value = f(seed, step, env, slot)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int r = 0; r < 10; ++r) {
    const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

__device__ __forceinline__ float u32_to_unit(uint32_t x) { return (float)(x >> 8) * 5.9604644775390625e-08f; }  // [0,1)

struct Rng {
  uint64_t seed, step;
  // uniform [0,1) for (env, slot)
  __device__ __forceinline__ float uniform(int env, int slot) const {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)env, (uint32_t)(slot >> 2), (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const int c = slot & 3;
    return u32_to_unit(c == 0 ? r.x : c == 1 ? r.y : c == 2 ? r.z : r.w);
  }
  // all four uniforms of one counter block
  __device__ __forceinline__ float4 uniform4(int env, int block) const {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)env, (uint32_t)block, (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    return make_float4(u32_to_unit(r.x), u32_to_unit(r.y), u32_to_unit(r.z), u32_to_unit(r.w));
  }
  // two uniforms of one counter block (used with block >= 64: disjoint from the slot blocks above)
  __device__ __forceinline__ float2 uniform2(int env, int block) const {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)env, (uint32_t)block, (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    return make_float2(u32_to_unit(r.x), u32_to_unit(r.y));
  }
  // four standard normals from ONE counter block (two Box-Muller pairs); `block` must be unique per (env, use)
  __device__ __forceinline__ float4 normal4(int env, uint32_t block) const {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)env, 0x40000000u | block, (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const float two_pi = 6.2831854820251465f;
    const float r0 = sqrtf(-2.0f * __logf(1.0f - u32_to_unit(r.x))), r1 = sqrtf(-2.0f * __logf(1.0f - u32_to_unit(r.z)));
    float s0, c0, s1, c1;
    __sincosf(two_pi * u32_to_unit(r.y), &s0, &c0);
    __sincosf(two_pi * u32_to_unit(r.w), &s1, &c1);
    return make_float4(r0 * c0, r0 * s0, r1 * c1, r1 * s1);
  }
  // three standard normals of one env from ONE counter block (the kicked object's force components)
  __device__ __forceinline__ void normal3(int env, uint32_t block, float (&o)[3]) const {
    const float4 g = normal4(env, 0x10000000u | block);
    o[0] = g.x; o[1] = g.y; o[2] = g.z;
  }
  // standard normal for (env, slot): Box-Muller over an independent counter block
  __device__ __forceinline__ float normal(int env, int slot) const {
    const uint4 r = philox4x32_10(make_uint4((uint32_t)env, 0x80000000u | (uint32_t)(slot >> 1), (uint32_t)step, (uint32_t)(step >> 32)),
                                  make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
    const float u1 = 1.0f - u32_to_unit((slot & 1) ? r.z : r.x);  // (0,1]
    const float u2 = u32_to_unit((slot & 1) ? r.w : r.y);
    return sqrtf(-2.0f * logf(u1)) * cosf(6.2831854820251465f * u2);
  }
};

// ---------------------------------------------------------------------------
// warp / block reductions in a fixed order (deterministic)
// ---------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
// Warp reductions on the redux unit (one instruction per 32-bit word instead of five dependent shuffle + op rounds: at the end of a
// persistent CTA these sit on the launch's critical path).  Maximum of floats through an order-preserving unsigned key (no NaNs here).
__device__ __forceinline__ float warp_max(float v) {
  const unsigned b = __float_as_uint(v);
  const unsigned k = __reduce_max_sync(0xffffffffu, (b & 0x80000000u) ? ~b : (b | 0x80000000u));
  return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

__device__ __forceinline__ float ldg(const float* p) { return __ldg(p); }

// Programmatic dependent launch (griddepcontrol): a kernel launched with the programmatic-stream-serialization
// attribute starts while its predecessor drains; nothing may touch global memory before pdl_wait().
// Both are no-ops when the kernel was launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// ---- in-kernel timeline (diagnostic builds only: -DHA_TRACE; see tools/trace_kernels.py) ----
// Thread 0 of every CTA records the SM clock at named points and %globaltimer at CTA start / end into a device array that
// ha_debug_trace_read copies out.  Compiles to nothing in the product build.
#ifdef HA_TRACE
struct HaTraceRec { unsigned long long gt0, gt1; long long c[14]; int smid, n; };
constexpr int kTraceMaxCtas = 4096;
static __device__ HaTraceRec g_ha_trace[2][kTraceMaxCtas];     // [0] pre kernel, [1] post kernel (one copy per translation unit; ha_kuka.cu's is read)
__device__ __forceinline__ unsigned long long ha_globaltimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ int ha_smid() { int s; asm volatile("mov.u32 %0, %%smid;" : "=r"(s)); return s; }
#define HA_TRACE_BEGIN(k) do { if (threadIdx.x == 0 && blockIdx.x < ha::kTraceMaxCtas) { ha::HaTraceRec& r_ = ha::g_ha_trace[k][blockIdx.x]; \
  r_.gt0 = ha::ha_globaltimer(); r_.gt1 = 0; r_.smid = ha::ha_smid(); r_.n = 1; _Pragma("unroll") for (int i_ = 1; i_ < 14; ++i_) r_.c[i_] = 0; r_.c[0] = clock64(); } } while (0)
#define HA_TRACE_POINT(k, i) do { if (threadIdx.x == 0 && blockIdx.x < ha::kTraceMaxCtas && (i) < 14) ha::g_ha_trace[k][blockIdx.x].c[i] = clock64(); } while (0)   /* stores only: nothing the warp waits for */
#define HA_TRACE_POINT_IF(k, i, cond) do { if ((cond) && blockIdx.x < ha::kTraceMaxCtas && (i) < 14) ha::g_ha_trace[k][blockIdx.x].c[i] = clock64(); } while (0)
/* CTA end = the LAST warp's end (every warp's lane 0 takes part) */
#define HA_TRACE_END(k) do { if ((threadIdx.x & 31) == 0 && blockIdx.x < ha::kTraceMaxCtas) atomicMax(&ha::g_ha_trace[k][blockIdx.x].gt1, ha::ha_globaltimer()); } while (0)
#else
#define HA_TRACE_BEGIN(k) do { } while (0)
#define HA_TRACE_POINT(k, i) do { } while (0)
#define HA_TRACE_POINT_IF(k, i, cond) do { } while (0)
#define HA_TRACE_END(k) do { } while (0)
#endif

// ---- cp.async (LDGSTS) ----
__device__ __forceinline__ void cp_async4(float* smem_dst, const float* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((unsigned)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }


// 64-bit sum (two's complement, exact mod 2^64): four 16-bit limbs, each limb sum of 32 lanes fits 21 bits
__device__ __forceinline__ long long warp_sum_ll(long long v) {
  const unsigned long long u = (unsigned long long)v;
  const unsigned long long s0 = __reduce_add_sync(0xffffffffu, (unsigned)(u & 0xffffu));
  const unsigned long long s1 = __reduce_add_sync(0xffffffffu, (unsigned)((u >> 16) & 0xffffu));
  const unsigned long long s2 = __reduce_add_sync(0xffffffffu, (unsigned)((u >> 32) & 0xffffu));
  const unsigned long long s3 = __reduce_add_sync(0xffffffffu, (unsigned)(u >> 48));
  return (long long)(s0 + (s1 << 16) + (s2 << 32) + (s3 << 48));
}
__device__ __forceinline__ int warp_sum_i(int v) { return __reduce_add_sync(0xffffffffu, v); }


// a / b with r = RN(1 / b) precomputed on the host: one Newton-style residual step (Markstein) gives the
// correctly rounded quotient for all but a vanishing fraction of operands (<= 1 ulp otherwise), in 3 instructions
__device__ __forceinline__ float div_by_const(float a, float b, float r) {
  const float q = a * r;
  const float e = fmaf(-b, q, a);
  return fmaf(e, r, q);
}


// order-preserving float <-> int64 key (for atomicMax on signed integers)
__device__ __forceinline__ long long float_key(float f) {
  const int b = __float_as_int(f);
  return (long long)(b >= 0 ? b : (b ^ 0x7fffffff));
}
__device__ __forceinline__ float key_float(long long k) {
  const int b = (int)k;
  return __int_as_float(b >= 0 ? b : (b ^ 0x7fffffff));
}

// ---- bulk asynchronous copies (TMA engine, non-tensor form: SASS UBLKCP) + mbarrier transaction tracking ----
__device__ __forceinline__ unsigned smem_u32(const void* p) { return (unsigned)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(unsigned long long* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive_expect_tx(unsigned long long* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "MBAR_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra MBAR_DONE_%=;\n"
      "bra MBAR_WAIT_%=;\n"
      "MBAR_DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
// global -> shared, `bytes` a multiple of 16, both addresses 16-byte aligned; completion is counted on `bar`
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, unsigned bytes, unsigned long long* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)), "l"(gsrc),
               "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
// shared -> global (bulk group of the issuing thread)
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read0() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }   // sources may be overwritten
__device__ __forceinline__ void bulk_wait0() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
// generic-proxy writes to shared memory become visible to the async proxy (bulk copies reading that memory)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// column descriptor of the observation-row assembly (hand and DeXtreme post kernels)
enum { kOpCopy = 0, kOpScale = 1, kOpUnscale = 2 };
struct HandCol { int base, stride, op; float a, b, c; };   // value(e) = stage[base + e * stride]; scale: * a; unscale: lo = a, hi = b, 1/(hi-lo) = c

// Row assembly, warp per env row: lane l owns columns l, l + 32, ... and keeps their descriptors in registers, so one row costs
// ceil(W / 32) shared loads, ops and 128-byte coalesced stores per warp.  `rank` / `nwarps`: this warp's index among the warps taking part.
// `map` (nullable): column j of the output is column map[j] of the descriptor table (a permuted / subset view of the row).
template <int MAX_PER_LANE>
__device__ __forceinline__ void assemble_rows(const float* stage, const HandCol* cols, float* __restrict__ out, const int W, const int nvalid,
                                              const float clip, const int lane, const int rank, const int nwarps, const int* map = nullptr) {
  HandCol c[MAX_PER_LANE];
#pragma unroll
  for (int k = 0; k < MAX_PER_LANE; ++k) {
    const int j = lane + 32 * k;
    c[k] = (j < W) ? cols[map ? __ldg(map + j) : j] : HandCol{0, 0, kOpCopy, 0.f, 0.f, 0.f};
  }
  const bool do_clip = clip > 0.0f;
  for (int e = rank; e < nvalid; e += nwarps) {
    float* row = out + (size_t)e * W;
#pragma unroll
    for (int k = 0; k < MAX_PER_LANE; ++k) {
      const int j = lane + 32 * k;
      if (j < W) {
        float v = stage[c[k].base + e * c[k].stride];
        if (c[k].op == kOpScale) v = c[k].a * v;
        else if (c[k].op == kOpUnscale) v = div_by_const(2.0f * v - c[k].b - c[k].a, c[k].b - c[k].a, c[k].c);   // torch_jit_utils.py:238-240
        if (do_clip) v = fminf(fmaxf(v, -clip), clip);
        row[j] = v;
      }
    }
  }
}

}  // namespace ha
