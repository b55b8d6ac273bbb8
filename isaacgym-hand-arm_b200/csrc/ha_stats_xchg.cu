// Cross-rank episode statistics over NVLink peer memory; see include/handarm_b200.h (HaStatsExchange).
// One single-CTA launch per exchange: push this rank's 16 floats + a sequence tag into every peer's buffer with plain remote stores,
// and reduce the newest exchange whose vectors have all arrived.  No collective kernel next to the persistent step kernels, no waiting.
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

// exchange buffer of one rank: float slots[RING][world][HA_STATS_LEN], long long tags[RING][world], long long head[world]
__host__ __device__ inline size_t xchg_tag_offset_floats(int world) { return (size_t)HA_XCHG_RING * world * HA_STATS_LEN; }

__device__ __forceinline__ void st_release_sys(long long* p, long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ void st_relaxed_sys(long long* p, long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long ld_acquire_sys(const long long* p) {
  long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

constexpr int kXchgPushThreads = HA_XCHG_MAX_RANKS * HA_STATS_LEN;      // thread = (peer, statistic)
constexpr int kXchgThreads = kXchgPushThreads + 32;                       // + one warp that reduces meanwhile

// Wait-free on both sides.  Writer (exchange number seq, slot seq % RING of every peer): tag = -seq (slot under construction),
// fence, the 16 values, tag = seq, head[rank] = seq.  Reader (the spare warp): m = min over the ranks of head[r] = the newest
// exchange every rank has published; if it is newer than the last one reduced and not about to be overwritten, read the W
// vectors of slot m % RING between two looks at their tags (seqlock) and reduce them in rank order; otherwise `reduced` keeps
// its value.  Nobody ever waits: a rank that is late only delays the exchange number the others reduce.
__global__ void __launch_bounds__(kXchgThreads)
stats_exchange_kernel(const HaStatsExchange x, const float* __restrict__ stats, float* __restrict__ reduced) {
  const int tid = threadIdx.x, W = x.world;
  const long long seq = x.seq[0] + 1;                   // every thread reads the same value; thread 0 publishes it at the end
  const int ring = (int)(seq % HA_XCHG_RING);
  const size_t tag_ofs = xchg_tag_offset_floats(W);
  if (tid < kXchgPushThreads) {
    const int peer = tid / HA_STATS_LEN, i = tid % HA_STATS_LEN;
    if (peer < W && i == 0) {
      long long* tags = reinterpret_cast<long long*>(x.peer_bufs[peer] + tag_ofs);
      st_relaxed_sys(tags + (size_t)ring * W + x.rank, -seq);
      __threadfence_system();                           // the slot is marked before any of its values changes
    }
  } else {
    // ---- the spare warp: reduce the newest exchange that is complete on this rank ----
    const int lane = tid - kXchgPushThreads;
    float* mine = x.peer_bufs[x.rank];
    const long long* tags = reinterpret_cast<const long long*>(mine + tag_ofs);
    const long long* head = tags + (size_t)HA_XCHG_RING * W;
    long long h = 0x7fffffffffffffffLL;
    if (lane < W) h = ld_acquire_sys(head + lane);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) { const long long t = __shfl_xor_sync(0xffffffffu, h, o); h = t < h ? t : h; }
    const long long m = h, last = x.seq[1];
    bool ok = m >= 1 && m > last && seq - m < HA_XCHG_RING - 1;
    const int pr = (int)((m > 0 ? m : 0) % HA_XCHG_RING);
    bool tag_ok = true;
    if (ok && lane < W) tag_ok = ld_acquire_sys(tags + (size_t)pr * W + lane) == m;
    ok = ok && __all_sync(0xffffffffu, tag_ok);
    float acc = 0.f;
    if (ok && lane < HA_STATS_LEN) {
      const bool is_max = (x.max_mask >> lane) & 1u;
      const volatile float* slot = mine + (size_t)pr * W * HA_STATS_LEN + lane;
      acc = slot[0];
      for (int r = 1; r < W; ++r) {                      // rank order: the same association on every rank
        const float v = slot[(size_t)r * HA_STATS_LEN];
        acc = is_max ? fmaxf(acc, v) : acc + v;
      }
    }
    __threadfence_system();
    if (ok && lane < W) tag_ok = ld_acquire_sys(tags + (size_t)pr * W + lane) == m;      // nobody began to overwrite the slot meanwhile
    ok = ok && __all_sync(0xffffffffu, tag_ok);
    if (ok && lane < HA_STATS_LEN) reduced[lane] = acc;
    if (lane == 0) {
      if (ok) x.seq[1] = m;
      else if (m >= 1 && m > last && seq - m >= HA_XCHG_RING - 1) *x.status = 2;       // a peer is a whole ring behind
    }
  }
  __syncthreads();
  if (tid < kXchgPushThreads) {
    const int peer = tid / HA_STATS_LEN, i = tid % HA_STATS_LEN;
    if (peer < W) x.peer_bufs[peer][((size_t)ring * W + x.rank) * HA_STATS_LEN + i] = stats[i];
  }
  __syncthreads();       // the values of all threads happen-before the tags below (release is cumulative over the barrier)
  if (tid < W) {
    long long* tags = reinterpret_cast<long long*>(x.peer_bufs[tid] + tag_ofs);
    st_release_sys(tags + (size_t)ring * W + x.rank, seq);
    st_release_sys(tags + (size_t)HA_XCHG_RING * W + x.rank, seq);        // head[rank]
  }
  if (tid == 0) x.seq[0] = seq;
}

}  // namespace ha

extern "C" {

size_t ha_sizeof_stats_exchange(void) { return sizeof(HaStatsExchange); }

size_t ha_stats_exchange_bytes(int32_t world) {
  if (world < 1 || world > HA_XCHG_MAX_RANKS) return 0;
  return ha::xchg_tag_offset_floats(world) * sizeof(float) + ((size_t)HA_XCHG_RING + 1) * world * sizeof(long long);
}

int ha_stats_exchange(const HaStatsExchange* x, const float* stats, float* reduced, ha_stream_t stream) {
  HA_REQUIRE(x != nullptr && stats != nullptr && reduced != nullptr, "null argument");
  HA_REQUIRE(x->world >= 1 && x->world <= HA_XCHG_MAX_RANKS, "world must be in 1..%d (got %d)", HA_XCHG_MAX_RANKS, x->world);
  HA_REQUIRE(x->rank >= 0 && x->rank < x->world, "rank %d outside of world %d", x->rank, x->world);
  HA_REQUIRE(x->peer_bufs && x->seq && x->status, "HaStatsExchange has a null field");
  return ha::check_cuda(ha::launch(ha::stats_exchange_kernel, 1, ha::kXchgThreads, 0, (cudaStream_t)stream, *x, stats, reduced),
                        "stats_exchange_kernel launch");
}

}  // extern "C"
