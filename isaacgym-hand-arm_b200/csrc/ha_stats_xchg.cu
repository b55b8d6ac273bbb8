// Cross-rank episode statistics over NVLink peer memory; see include/handarm_b200.h (HaStatsExchange).
// One single-warp launch per exchange: push this rank's 16 floats + a sequence tag into every peer's buffer with plain remote stores,
// and reduce the newest exchange whose vectors have all arrived.  No collective kernel next to the persistent step kernels, no waiting.
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

#include "ha_stats_xchg.cuh"

namespace ha {

// the stand-alone launch: one warp (the post-physics kernel runs the same function in one of its own warps, see HaKukaBuffers.xchg)
__global__ void __launch_bounds__(32)
stats_exchange_kernel(const HaStatsExchange x, const float* __restrict__ stats, float* __restrict__ reduced) {
  stats_exchange_warp(x, stats, reduced, (int)threadIdx.x);
}

}  // namespace ha

extern "C" {

size_t ha_sizeof_stats_exchange(void) { return sizeof(HaStatsExchange); }

size_t ha_stats_exchange_bytes(int32_t world) {
  if (world < 1 || world > HA_XCHG_MAX_RANKS) return 0;
  return (ha::xchg_words(world) + (size_t)world) * sizeof(unsigned long long);      // tagged words + head[world]
}

int ha_stats_exchange(const HaStatsExchange* x, const float* stats, float* reduced, ha_stream_t stream) {
  HA_REQUIRE(x != nullptr && stats != nullptr && reduced != nullptr, "null argument");
  HA_REQUIRE(x->world >= 1 && x->world <= HA_XCHG_MAX_RANKS, "world must be in 1..%d (got %d)", HA_XCHG_MAX_RANKS, x->world);
  HA_REQUIRE(x->rank >= 0 && x->rank < x->world, "rank %d outside of world %d", x->rank, x->world);
  HA_REQUIRE(x->peer_bufs && x->seq && x->status, "HaStatsExchange has a null field");
  HA_REQUIRE(ha::aligned8(stats) && ha::aligned8(reduced), "stats / reduced must be 8-byte aligned");
  return ha::check_cuda(ha::launch(ha::stats_exchange_kernel, 1, 32, 0, (cudaStream_t)stream, *x, stats, reduced),
                        "stats_exchange_kernel launch");
}

}  // extern "C"
