// Cross-rank episode statistics over NVLink peer memory; see include/handarm_b200.h (HaStatsExchange).
// One single-CTA launch per exchange: push this rank's 16 floats + a sequence tag into every peer's buffer with plain remote stores,
// then reduce what the peers pushed during the previous exchange.  No collective kernel next to the persistent step kernels.
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

// exchange buffer of one rank: float slots[RING][world][HA_STATS_LEN], then long long tags[RING][world]
__host__ __device__ inline size_t xchg_tag_offset_floats(int world) { return (size_t)HA_XCHG_RING * world * HA_STATS_LEN; }

__device__ __forceinline__ void st_release_sys(long long* p, long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ long long ld_acquire_sys(const long long* p) {
  long long v;
  asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

constexpr int kXchgPushThreads = HA_XCHG_MAX_RANKS * HA_STATS_LEN;      // thread = (peer, statistic)
constexpr int kXchgThreads = kXchgPushThreads + 32;                       // + one warp that reduces the previous exchange meanwhile

__global__ void __launch_bounds__(kXchgThreads)
stats_exchange_kernel(const HaStatsExchange x, const float* __restrict__ stats, float* __restrict__ reduced) {
  const int tid = threadIdx.x, W = x.world;
  const long long seq = *x.seq + 1;                     // every thread reads the same value; thread 0 publishes it at the end
  const int ring = (int)(seq % HA_XCHG_RING);
  if (tid < kXchgPushThreads) {
    // ---- push: thread (peer, i) stores statistic i into slot [ring][rank] of that peer's buffer ----
    const int peer = tid / HA_STATS_LEN, i = tid % HA_STATS_LEN;
    if (peer < W) x.peer_bufs[peer][((size_t)ring * W + x.rank) * HA_STATS_LEN + i] = stats[i];
  } else if (seq > 1) {
    // ---- the last warp reduces what arrived for the previous exchange (a whole step loop ago: normally no wait at all) ----
    const int lane = tid - kXchgPushThreads;
    const long long prev = seq - 1;
    const int pr = (int)(prev % HA_XCHG_RING);
    float* mine = x.peer_bufs[x.rank];
    if (lane < W) {
      const long long* tag = reinterpret_cast<const long long*>(mine + xchg_tag_offset_floats(W)) + (size_t)pr * W + lane;
      int polls = 0;
      while (ld_acquire_sys(tag) != prev) {
        if (++polls > HA_XCHG_SPIN_LIMIT) { *x.status = 1; break; }
        __nanosleep(200);
      }
    }
    __syncwarp();
    if (lane < HA_STATS_LEN) {
      const bool is_max = (x.max_mask >> lane) & 1u;
      const volatile float* slot = mine + (size_t)pr * W * HA_STATS_LEN + lane;
      float acc = slot[0];
      for (int r = 1; r < W; ++r) {                      // rank order: the same association on every rank
        const float v = slot[(size_t)r * HA_STATS_LEN];
        acc = is_max ? fmaxf(acc, v) : acc + v;
      }
      reduced[lane] = acc;
    }
  }
  __syncthreads();       // the pushes of all threads happen-before the tags below (release is cumulative over the barrier)
  if (tid < W) {
    long long* tags = reinterpret_cast<long long*>(x.peer_bufs[tid] + xchg_tag_offset_floats(W));
    st_release_sys(tags + (size_t)ring * W + x.rank, seq);
  }
  if (tid == 0) *x.seq = seq;
}

}  // namespace ha

extern "C" {

size_t ha_sizeof_stats_exchange(void) { return sizeof(HaStatsExchange); }

size_t ha_stats_exchange_bytes(int32_t world) {
  if (world < 1 || world > HA_XCHG_MAX_RANKS) return 0;
  return ha::xchg_tag_offset_floats(world) * sizeof(float) + (size_t)HA_XCHG_RING * world * sizeof(long long);
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
