// Tensor-core layers of the random network adversary (rna_util.py:128-136): D[M,N] = epilogue(A[M,K] . W[N,K]^T) on the 5th-gen
// tensor cores (tcgen05.mma, accumulators in TMEM, operands staged by TMA), one launch per wide layer.
//
// fp32 fidelity on bf16 tensor cores: every fp32 operand x is carried as two bf16 planes, hi = bf16(x) and lo = bf16(x - hi)
// (16 mantissa bits together), and a k-block is accumulated three times into the same fp32 TMEM tile:
//     A_hi.W_hi + A_hi.W_lo + A_lo.W_hi          (A_lo.W_lo, ~2^-16 relative, is dropped)
// so the logits agree with an fp32 GEMM to ~1e-5 -- the argmax the adversary takes is the reference's wherever the top-2 margin
// exceeds that (tests/test_rna_gpu.py) -- at 3 bf16 MMAs per product instead of an fp32 SIMT GEMM (51 TFLOP/s measured for the
// library fp32 path on B200).
//
// CTA = one 128 x 128 output tile, 6 warps: warp 0 lane 0 = TMA producer (3 stages; a stage is one k-block of 32 with all four
// operand tiles A_hi, A_lo, W_hi, W_lo of 128 x 32 bf16 in 64-byte swizzled rows, so each tile is fetched once for the three
// products: 96 flop per L2 byte instead of 64 -- the first version, one (A, W) pair per stage, ran at the L2 -> SM bandwidth),
// warp 1 = TMEM allocation, lane 0 = MMA issuer (3 products x 2 tcgen05.mma K=16 per stage, tcgen05.commit frees the stage), warps 2..5 =
// epilogue: tcgen05.ld 32 columns at a time (thread = output row), then either
//   kEpiAct     + bias, ReLU, * dropout mask, split to bf16 hi / lo planes (the next layer's operand), 64-byte stores
//   kEpiSelect  + bias, argmax over the 32 bins of a joint (one tcgen05.ld chunk = one joint), gather from discretised_dofs
// Two CTAs fit an SM (97 KB smem, 128 TMEM columns each), so one tile's epilogue overlaps the other's main loop.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdlib.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

constexpr int kBM = 128, kBN = 128, kBK = 32, kStages = 3, kUmmaK = 16;
constexpr int kTcThreads = 192;
constexpr unsigned kTileBytes = kBM * kBK * 2;                        // 8 KB: 128 rows of 64 bytes
constexpr unsigned kStageBytes = 4 * kTileBytes;                       // A_hi, A_lo, W_hi, W_lo
static_assert(kBM == kBN, "operand tiles share one box shape");
constexpr unsigned kTmemCols = 128;
enum { kEpiAct = 0, kEpiSelect = 1 };

struct RnaTcSmem {                       // after the 1024-byte aligned stage area
  unsigned long long full[kStages], empty[kStages], tmem_full;
  unsigned tmem_base;
};
constexpr size_t kTcSmemBytes = 1024 + (size_t)kStages * kStageBytes + sizeof(RnaTcSmem);

__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, unsigned long long* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(smem_dst)),
               "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_commit(unsigned long long* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem desc] . B[smem desc]^T, bf16 operands, fp32 accumulate
__device__ __forceinline__ void tc_mma_bf16(unsigned d_tmem, unsigned long long a_desc, unsigned long long b_desc, unsigned idesc, unsigned accumulate) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "setp.ne.b32 p, %4, 0;\n"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
      "}\n" ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
// shared-memory matrix descriptor: K-major tile of 64-byte rows, 64-byte swizzle (layout type 4), 8-row groups 512 bytes apart
__device__ __forceinline__ unsigned long long tc_smem_desc(const void* tile) {
  const unsigned long long addr = (unsigned long long)(smem_u32(tile) & 0x3FFFFu) >> 4;
  return addr | (1ull << 16) | ((unsigned long long)(512 >> 4) << 32) | (1ull << 46) | (4ull << 61);
}
// 32 lanes x 32 columns of fp32 accumulators: thread = lane (row), v[j] = column j
__device__ __forceinline__ void tc_ld32(unsigned taddr, float (&v)[32]) {
  unsigned r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]),
        "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]),
        "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]),
        "=r"(r[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}
__device__ __forceinline__ unsigned pack_bf16(__nv_bfloat16 a, __nv_bfloat16 b) {
  return (unsigned)__bfloat16_as_ushort(a) | ((unsigned)__bfloat16_as_ushort(b) << 16);
}

struct RnaTcArgs {
  int M, N, K;
  const float* bias;                 // [N]
  const uint32_t* mask_bits;         // kEpiAct: nullable [M, N / 32], bit j of word (row, c) = dropout mask of column 32 c + j
  int relu;
  __nv_bfloat16* out_hi;             // kEpiAct: [M, N] planes of the result
  __nv_bfloat16* out_lo;
  const float* discretised;          // kEpiSelect: [16, 32]
  float* out;                        // kEpiSelect: [M, 16]
  int32_t* idx_out;                  // kEpiSelect: nullable [M, 16]
};

// Epilogue of 32 consecutive columns [col0, col0 + 32) of one output row (v = fp32 accumulators of that row)
// `bias32`: the 32 bias values of these columns (global or shared memory, 16-byte aligned)
template <int EPI>
__device__ __forceinline__ void rna_epi_chunk(const RnaTcArgs& g, float (&v)[32], const int row, const int col0, const float* bias32) {
#pragma unroll
  for (int j = 0; j < 32; j += 4) {
    const float4 bv = *reinterpret_cast<const float4*>(bias32 + j);
    v[j] += bv.x; v[j + 1] += bv.y; v[j + 2] += bv.z; v[j + 3] += bv.w;
  }
  if (row < g.M) {
    if constexpr (EPI == kEpiAct) {
      if (g.relu) {
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = fmaxf(v[j], 0.f);
      }
      if (g.mask_bits) {                           // dropout_masks * x (the mask is {0, 1}): one word per 32 columns
        const unsigned m = __ldg(g.mask_bits + (size_t)row * (g.N >> 5) + (col0 >> 5));
#pragma unroll
        for (int j = 0; j < 32; ++j) v[j] = ((m >> j) & 1u) ? v[j] : 0.0f * v[j];
      }
      unsigned hi[16], lo[16];
#pragma unroll
      for (int j = 0; j < 32; j += 2) {          // packed conversions: one cvt.rn.bf16x2.f32 per pair; bf16 -> fp32 is a shift
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[j], v[j + 1]);
        const unsigned hb = *reinterpret_cast<const unsigned*>(&h);
        const __nv_bfloat162 l = __floats2bfloat162_rn(v[j] - __uint_as_float(hb << 16), v[j + 1] - __uint_as_float(hb & 0xffff0000u));
        hi[j >> 1] = hb;
        lo[j >> 1] = *reinterpret_cast<const unsigned*>(&l);
      }
      uint4* dh = reinterpret_cast<uint4*>(g.out_hi + (size_t)row * g.N + col0);
      uint4* dl = reinterpret_cast<uint4*>(g.out_lo + (size_t)row * g.N + col0);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        dh[j] = make_uint4(hi[4 * j], hi[4 * j + 1], hi[4 * j + 2], hi[4 * j + 3]);
        dl[j] = make_uint4(lo[4 * j], lo[4 * j + 1], lo[4 * j + 2], lo[4 * j + 3]);
      }
    } else {
      // softmax is monotone: bin = first maximum of the logits (torch.argmax), value = discretised_dofs[d, bin]
      const int d = col0 / HA_RNA_BINS;
      float best = v[0];
      int idx = 0;
#pragma unroll
      for (int j = 1; j < 32; ++j) {
        if (v[j] > best) { best = v[j]; idx = j; }
      }
      g.out[(size_t)row * HA_RNA_DOFS + d] = __ldg(g.discretised + d * HA_RNA_BINS + idx);
      if (g.idx_out) g.idx_out[(size_t)row * HA_RNA_DOFS + d] = idx;
    }
  }
}

template <int EPI>
__global__ void __launch_bounds__(kTcThreads, 2)
rna_gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                   const __grid_constant__ CUtensorMap map_w_hi, const __grid_constant__ CUtensorMap map_w_lo,
                   const __grid_constant__ RnaTcArgs g) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* stages = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  RnaTcSmem& sm = *reinterpret_cast<RnaTcSmem*>(stages + (size_t)kStages * kStageBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * kBN, m0 = blockIdx.y * kBM;
  const int n_iters = g.K / kBK;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kStages; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], 1); }
    mbar_init(&sm.tmem_full, 1);
    mbar_fence_init();
  }
  if (warp == 1) {                                   // TMEM: 128 fp32 columns x 128 lanes for the tile's accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "n"(kTmemCols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: the four operand tiles of a k-block =====
      for (int it = 0; it < n_iters; ++it) {
        const int s = it % kStages;
        mbar_wait(&sm.empty[s], ((it / kStages) & 1) ^ 1);
        mbar_arrive_expect_tx(&sm.full[s], kStageBytes);
        unsigned char* dst = stages + (size_t)s * kStageBytes;
        tma_load_2d(dst, &map_a_hi, it * kBK, m0, &sm.full[s]);
        tma_load_2d(dst + kTileBytes, &map_a_lo, it * kBK, m0, &sm.full[s]);
        tma_load_2d(dst + 2 * kTileBytes, &map_w_hi, it * kBK, n0, &sm.full[s]);
        tma_load_2d(dst + 3 * kTileBytes, &map_w_lo, it * kBK, n0, &sm.full[s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      // instruction descriptor: D = F32, A = B = BF16, both K-major, N = 128, M = 128
      constexpr unsigned idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(kBN >> 3) << 17) | ((unsigned)(kBM >> 4) << 24);
      for (int it = 0; it < n_iters; ++it) {
        const int s = it % kStages;
        mbar_wait(&sm.full[s], (it / kStages) & 1);
        tc_fence_after();
        const unsigned char* src = stages + (size_t)s * kStageBytes;
        const unsigned long long a_hi = tc_smem_desc(src), a_lo = tc_smem_desc(src + kTileBytes);
        const unsigned long long w_hi = tc_smem_desc(src + 2 * kTileBytes), w_lo = tc_smem_desc(src + 3 * kTileBytes);
#pragma unroll
        for (int k = 0; k < kBK / kUmmaK; ++k) {   // 32 bytes further along K inside the swizzle atom per K = 16 slice
          tc_mma_bf16(tmem, a_hi + 2ull * k, w_hi + 2ull * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
          tc_mma_bf16(tmem, a_hi + 2ull * k, w_lo + 2ull * k, idesc, 1u);
          tc_mma_bf16(tmem, a_lo + 2ull * k, w_hi + 2ull * k, idesc, 1u);
        }
        tc_commit(&sm.empty[s]);                    // the stage is free once these MMAs have read it
      }
      tc_commit(&sm.tmem_full);                     // accumulators complete
    }
  } else {
    // ===== epilogue: warp q owns TMEM lanes 32q .. 32q + 31 (q = warp % 4), thread = output row =====
    const int q = warp & 3;
    const int row = m0 + 32 * q + lane;
    mbar_wait(&sm.tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c = 0; c < kBN / 32; ++c) {
      float v[32];
      tc_ld32(tmem + ((unsigned)(32 * q) << 16) + (unsigned)(32 * c), v);
      const int col0 = n0 + 32 * c;
      rna_epi_chunk<EPI>(g, v, row, col0, g.bias + col0);
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(kTmemCols) : "memory");
  }
}

// ---------------------------------------------------------------------------------------------------------------------------
// Persistent version (the default): one CTA per SM walks 128 x 256 output tiles (tile = blockIdx.x, + gridDim.x, ...; the N tiles of
// a row block are neighbours, so concurrent CTAs share their A rows in L2).  What it removes from the one-tile-per-CTA kernel
// above, measured at 8-12 us of fixed cost per 24-37 us tile: barrier / TMEM set-up once per CTA instead of per tile, a TMA ring
// that never drains between tiles (stage = A_hi, A_lo 128 x kPBK and W_hi, W_lo 256 x kPBK), and two accumulator buffers
// in TMEM (2 x 256 columns) so the epilogue of tile i runs under the main loop of tile i + 1.  N = 256 per MMA halves the A traffic
// per flop (128 flop per L2 byte): the L2 -> SM path (~12 TB/s chip-wide) is what bounded the first version's main loop.
#ifndef HA_RNA_PBK
#define HA_RNA_PBK 64      // k-block of the persistent kernel: 64 (128-byte swizzled rows, 2 stages of 96 KB) or 32 (64-byte rows, 4 stages of 48 KB)
#endif
constexpr int kPBN = 256, kPBK = HA_RNA_PBK, kPStages = (kPBK == 64) ? 2 : 4;
constexpr unsigned kPATile = kBM * kPBK * 2, kPWTile = kPBN * kPBK * 2;
constexpr unsigned kPStageBytes = 2 * kPATile + 2 * kPWTile;
#ifndef HA_RNA_EPI_WARPS
#define HA_RNA_EPI_WARPS 4    // epilogue warps of the persistent kernel (one per TMEM lane quarter); 16 (four per quarter) measured slower: 0.267 vs 0.253 ms
#endif
constexpr int kPEpiWarps = HA_RNA_EPI_WARPS, kPThreads = 64 + 32 * kPEpiWarps;
static_assert(kPEpiWarps % 4 == 0 && (kPBN / 32) % (kPEpiWarps / 4) == 0, "epilogue warps split the column chunks of their lane quarter evenly");
// K-major tile descriptor for the persistent kernel's row width: 128-byte rows / 128-byte swizzle (layout 2, 8-row groups 1024 bytes apart)
__device__ __forceinline__ unsigned long long tc_smem_desc_p(const void* tile) {
  if constexpr (kPBK == 32) return tc_smem_desc(tile);
  const unsigned long long addr = (unsigned long long)(smem_u32(tile) & 0x3FFFFu) >> 4;
  return addr | (1ull << 16) | ((unsigned long long)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
struct RnaTcPSmem {
  unsigned long long full[kPStages], empty[kPStages], tmem_full[2], tmem_empty[2];
  unsigned tmem_base;
  alignas(16) float bias[2][kPBN];     // the tile's bias slice per accumulator buffer, staged by the epilogue warps ahead of the wait
};
constexpr size_t kTcPSmemBytes = 1024 + (size_t)kPStages * kPStageBytes + sizeof(RnaTcPSmem);

template <int EPI>
__global__ void __launch_bounds__(kPThreads, 1)
rna_gemm_tc_persistent_kernel(const __grid_constant__ CUtensorMap map_a_hi, const __grid_constant__ CUtensorMap map_a_lo,
                              const __grid_constant__ CUtensorMap map_w_hi, const __grid_constant__ CUtensorMap map_w_lo,
                              const __grid_constant__ RnaTcArgs g) {
  extern __shared__ unsigned char smem_raw[];
  unsigned char* stages = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  RnaTcPSmem& sm = *reinterpret_cast<RnaTcPSmem*>(stages + (size_t)kPStages * kPStageBytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_iters = g.K / kPBK;
  const int tiles_n = g.N / kPBN, tiles_m = (g.M + kBM - 1) / kBM, n_tiles = tiles_n * tiles_m;

  if (threadIdx.x == 0) {
#pragma unroll
    for (int s = 0; s < kPStages; ++s) { mbar_init(&sm.full[s], 1); mbar_init(&sm.empty[s], 1); }
#pragma unroll
    for (int b = 0; b < 2; ++b) { mbar_init(&sm.tmem_full[b], 1); mbar_init(&sm.tmem_empty[b], 32 * kPEpiWarps); }
    mbar_fence_init();
  }
  if (warp == 1) {                                   // all 512 TMEM columns: two 128 x 256 fp32 accumulator buffers
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&sm.tmem_base)), "n"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const unsigned tmem = sm.tmem_base;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: runs ahead across tiles, bounded only by the ring =====
      int git = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
        const int m0 = (tile / tiles_n) * kBM, n0 = (tile % tiles_n) * kPBN;
        for (int it = 0; it < n_iters; ++it, ++git) {
          const int s = git % kPStages;
          mbar_wait(&sm.empty[s], ((git / kPStages) & 1) ^ 1);
          mbar_arrive_expect_tx(&sm.full[s], kPStageBytes);
          unsigned char* dst = stages + (size_t)s * kPStageBytes;
          tma_load_2d(dst, &map_a_hi, it * kPBK, m0, &sm.full[s]);
          tma_load_2d(dst + kPATile, &map_a_lo, it * kPBK, m0, &sm.full[s]);
          tma_load_2d(dst + 2 * kPATile, &map_w_hi, it * kPBK, n0, &sm.full[s]);
          tma_load_2d(dst + 2 * kPATile + kPWTile, &map_w_lo, it * kPBK, n0, &sm.full[s]);
        }
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer: D = F32, A = B = BF16, K-major, M = 128, N = 256 =====
      constexpr unsigned idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((unsigned)(kPBN >> 3) << 17) | ((unsigned)(kBM >> 4) << 24);
      int git = 0, lt = 0;
      for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
        const int buf = lt & 1;
        mbar_wait(&sm.tmem_empty[buf], ((lt >> 1) & 1) ^ 1);          // the epilogue has drained this accumulator buffer
        tc_fence_after();
        const unsigned d_tmem = tmem + (unsigned)(buf * kPBN);
        for (int it = 0; it < n_iters; ++it, ++git) {
          const int s = git % kPStages;
          mbar_wait(&sm.full[s], (git / kPStages) & 1);
          tc_fence_after();
          const unsigned char* src = stages + (size_t)s * kPStageBytes;
          const unsigned long long a_hi = tc_smem_desc_p(src), a_lo = tc_smem_desc_p(src + kPATile);
          const unsigned long long w_hi = tc_smem_desc_p(src + 2 * kPATile), w_lo = tc_smem_desc_p(src + 2 * kPATile + kPWTile);
#pragma unroll
          for (int k = 0; k < kPBK / kUmmaK; ++k) {
            tc_mma_bf16(d_tmem, a_hi + 2ull * k, w_hi + 2ull * k, idesc, (it > 0 || k > 0) ? 1u : 0u);
            tc_mma_bf16(d_tmem, a_hi + 2ull * k, w_lo + 2ull * k, idesc, 1u);
            tc_mma_bf16(d_tmem, a_lo + 2ull * k, w_hi + 2ull * k, idesc, 1u);
          }
          tc_commit(&sm.empty[s]);
        }
        tc_commit(&sm.tmem_full[buf]);
      }
    }
  } else {
    // ===== epilogue warps: thread = output row; the kPEpiWarps / 4 warps of a TMEM lane quarter (warp % 4) share its 8 column chunks =====
    const int q = warp & 3, sub = (warp - 2) >> 2;
    int lt = 0;
    for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x, ++lt) {
      const int buf = lt & 1;
      const int m0 = (tile / tiles_n) * kBM, n0 = (tile % tiles_n) * kPBN;
      const int row = m0 + 32 * q + lane;
      for (int i = threadIdx.x - 64; i < kPBN; i += 32 * kPEpiWarps) sm.bias[buf][i] = __ldg(g.bias + n0 + i);
      asm volatile("bar.sync 1, %0;" ::"n"(32 * kPEpiWarps) : "memory");      // epilogue warps only
      mbar_wait(&sm.tmem_full[buf], (lt >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int c = sub; c < kPBN / 32; c += kPEpiWarps / 4) {
        float v[32];
        tc_ld32(tmem + ((unsigned)(32 * q) << 16) + (unsigned)(buf * kPBN + 32 * c), v);
        rna_epi_chunk<EPI>(g, v, row, n0 + 32 * c, &sm.bias[buf][32 * c]);
      }
      tc_fence_before();
      asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&sm.tmem_empty[buf])) : "memory");   // every epilogue thread arrives: the buffer is free
    }
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(512) : "memory");
  }
}

// fp32 -> (hi, lo) bf16 planes; used for the weights (once per refresh)
__global__ void __launch_bounds__(256) rna_split_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ hi, __nv_bfloat16* __restrict__ lo,
                                                          const size_t count) {
  const size_t stride = (size_t)gridDim.x * blockDim.x;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < count; i += stride) {
    const float v = x[i];
    const __nv_bfloat16 h = __float2bfloat16_rn(v);
    hi[i] = h;
    lo[i] = __float2bfloat16_rn(v - __bfloat162float(h));
  }
}

}  // namespace ha

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// [rows, cols] row-major bf16, box = 32 columns (64 bytes, one swizzle row) x 128 rows
int make_map(CUtensorMap* map, const void* ptr, int rows, int cols, int box_rows = ha::kBM, int box_cols = ha::kBK) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return ha::set_error(HA_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};      // box_cols bf16 = one swizzle row (64 or 128 bytes)
  const cuuint32_t estr[2] = {1, 1};
  const CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        box_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return ha::set_error(HA_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d (rows %d, cols %d)", (int)r, rows, cols);
  return HA_OK;
}

template <int EPI>
int launch_tc_tiles(const ha::RnaTcArgs& g, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, cudaStream_t st) {
  HA_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "M, N, K must be positive");
  HA_REQUIRE(g.N % ha::kBN == 0 && g.K % ha::kBK == 0, "N must be a multiple of %d and K of %d (got N %d, K %d)", ha::kBN, ha::kBK, g.N, g.K);
  HA_REQUIRE(a_hi && a_lo && w_hi && w_lo && g.bias, "null operand");
  HA_REQUIRE(ha::aligned16(a_hi) && ha::aligned16(a_lo) && ha::aligned16(w_hi) && ha::aligned16(w_lo) && ha::aligned16(g.bias), "operands must be 16-byte aligned");
  CUtensorMap ma_hi, ma_lo, mw_hi, mw_lo;
  int rc;
  if ((rc = make_map(&ma_hi, a_hi, g.M, g.K))) return rc;
  if ((rc = make_map(&ma_lo, a_lo, g.M, g.K))) return rc;
  if ((rc = make_map(&mw_hi, w_hi, g.N, g.K))) return rc;
  if ((rc = make_map(&mw_lo, w_lo, g.N, g.K))) return rc;
  static ha::PerDeviceFlag flag;
  bool& configured = flag.here();
  if (!configured) {
    rc = ha::check_cuda(cudaFuncSetAttribute(ha::rna_gemm_tc_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ha::kTcSmemBytes),
                        "cudaFuncSetAttribute(rna_gemm_tc_kernel)");
    if (rc) return rc;
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(g.N / ha::kBN), (unsigned)((g.M + ha::kBM - 1) / ha::kBM));
  cfg.blockDim = dim3(ha::kTcThreads);
  cfg.dynamicSmemBytes = ha::kTcSmemBytes;
  cfg.stream = st;
  return ha::check_cuda(cudaLaunchKernelEx(&cfg, ha::rna_gemm_tc_kernel<EPI>, ma_hi, ma_lo, mw_hi, mw_lo, g), "rna_gemm_tc_kernel launch");
}


bool tc_persistent() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("HA_RNA_TC_TILES");     // tuning hook: 1 selects the one-tile-per-CTA kernel
    v = (e && e[0] == '1') ? 0 : 1;
  }
  return v == 1;
}

template <int EPI>
int launch_tc(const ha::RnaTcArgs& g, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, cudaStream_t st) {
  if (!tc_persistent() || g.N % ha::kPBN != 0 || g.K % ha::kPBK != 0) return launch_tc_tiles<EPI>(g, a_hi, a_lo, w_hi, w_lo, st);
  HA_REQUIRE(g.M > 0 && g.N > 0 && g.K > 0, "M, N, K must be positive");
  HA_REQUIRE(g.K % ha::kPBK == 0, "K must be a multiple of %d (got %d)", ha::kPBK, g.K);
  HA_REQUIRE(a_hi && a_lo && w_hi && w_lo && g.bias, "null operand");
  HA_REQUIRE(ha::aligned16(a_hi) && ha::aligned16(a_lo) && ha::aligned16(w_hi) && ha::aligned16(w_lo) && ha::aligned16(g.bias), "operands must be 16-byte aligned");
  CUtensorMap ma_hi, ma_lo, mw_hi, mw_lo;
  int rc;
  if ((rc = make_map(&ma_hi, a_hi, g.M, g.K, ha::kBM, ha::kPBK))) return rc;
  if ((rc = make_map(&ma_lo, a_lo, g.M, g.K, ha::kBM, ha::kPBK))) return rc;
  if ((rc = make_map(&mw_hi, w_hi, g.N, g.K, ha::kPBN, ha::kPBK))) return rc;
  if ((rc = make_map(&mw_lo, w_lo, g.N, g.K, ha::kPBN, ha::kPBK))) return rc;
  static ha::PerDeviceFlag flag;
  bool& configured = flag.here();
  const int sm_count = ha::device_sm_count();
  if (!configured) {
    rc = ha::check_cuda(cudaFuncSetAttribute(ha::rna_gemm_tc_persistent_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ha::kTcPSmemBytes),
                        "cudaFuncSetAttribute(rna_gemm_tc_persistent_kernel)");
    if (rc) return rc;
    configured = true;
  }
  const int n_tiles = (g.N / ha::kPBN) * ((g.M + ha::kBM - 1) / ha::kBM);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)(n_tiles < sm_count ? n_tiles : sm_count));
  cfg.blockDim = dim3(ha::kPThreads);
  cfg.dynamicSmemBytes = ha::kTcPSmemBytes;
  cfg.stream = st;
  return ha::check_cuda(cudaLaunchKernelEx(&cfg, ha::rna_gemm_tc_persistent_kernel<EPI>, ma_hi, ma_lo, mw_hi, mw_lo, g), "rna_gemm_tc_persistent_kernel launch");
}

}  // namespace

extern "C" {

int ha_rna_split_bf16(const float* x, void* hi, void* lo, size_t count, ha_stream_t stream) {
  HA_REQUIRE(x && hi && lo, "null pointer");
  if (count == 0) return HA_OK;
  const size_t want = (count + 255) / 256;
  const int grid = (int)(want < 148 * 8 ? want : 148 * 8);
  return ha::check_cuda(ha::launch(ha::rna_split_kernel, grid, 256, 0, (cudaStream_t)stream, x, reinterpret_cast<__nv_bfloat16*>(hi),
                                   reinterpret_cast<__nv_bfloat16*>(lo), count), "rna_split_kernel launch");
}

int ha_rna_layer_tc(int32_t m, int32_t n, int32_t k, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, const float* bias,
                    const uint32_t* mask_bits, int32_t relu, void* out_hi, void* out_lo, ha_stream_t stream) {
  HA_REQUIRE(out_hi && out_lo, "null output");
  HA_REQUIRE(ha::aligned16(out_hi) && ha::aligned16(out_lo), "outputs must be 16-byte aligned");
  ha::RnaTcArgs g = {};
  g.M = m; g.N = n; g.K = k; g.bias = bias; g.mask_bits = mask_bits; g.relu = relu;
  g.out_hi = reinterpret_cast<__nv_bfloat16*>(out_hi);
  g.out_lo = reinterpret_cast<__nv_bfloat16*>(out_lo);
  return launch_tc<ha::kEpiAct>(g, a_hi, a_lo, w_hi, w_lo, (cudaStream_t)stream);
}

int ha_rna_select_tc(int32_t m, int32_t k, const void* a_hi, const void* a_lo, const void* w_hi, const void* w_lo, const float* bias3,
                     const float* discretised_dofs, float* out, int32_t* idx_out, ha_stream_t stream) {
  HA_REQUIRE(discretised_dofs && out, "null pointer");
  ha::RnaTcArgs g = {};
  g.M = m; g.N = HA_RNA_DOFS * HA_RNA_BINS; g.K = k; g.bias = bias3; g.discretised = discretised_dofs; g.out = out; g.idx_out = idx_out;
  return launch_tc<ha::kEpiSelect>(g, a_hi, a_lo, w_hi, w_lo, (cudaStream_t)stream);
}

}  // extern "C"
