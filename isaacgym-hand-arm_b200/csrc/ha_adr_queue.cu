// ADR boundary-worker objective queues on the device (adr_vec_task.py:760-812, the data half of adr_update); see
// include/handarm_b200.h.  The reference walks the ADR parameters in Python and, per parameter and bound, builds two boolean
// masks, gathers the objectives of the finished boundary workers, copies them to the host (.cpu().numpy().tolist(): one sync per
// queue, ~90 per step) and extends a deque(maxlen = adr_queue_threshold_length).  Here: one compaction launch over the envs and
// one launch with a CTA per queue; the host reads [mean, length] of every queue in a single copy and keeps the scalar range logic.
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

constexpr int kAdrThreads = 256;

__device__ __forceinline__ long long adr_ld_volatile(const int64_t* p) { return *reinterpret_cast<const volatile long long*>(p); }

// Launch 1: the finished boundary workers (rand_env_mask & worker_types == ADR_BOUNDARY, adr_vec_task.py:785-792) in ascending
// env order -> list[0 .. total).  Same order-preserving grid scan as the observer kernel (per-CTA counts tagged with the epoch).
__global__ void __launch_bounds__(kAdrThreads)
adr_compact_kernel(const int n, const int per_cta, const int all_envs, const __grid_constant__ HaAdrQueueBuffers b) {
  __shared__ int warp_cnt[kAdrThreads / 32];
  __shared__ long long red[2][kAdrThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long epoch = adr_ld_volatile(b.state + 1) + 1;
  const int beg = blockIdx.x * per_cta, end = min(n, beg + per_cta);
  auto qualifies = [&](int env) {
    if (env >= end) return false;
    const bool rnd = all_envs || (b.reset_buf[env] != 0 && !(b.apply_reset_buf && b.apply_reset_buf[env] != 0));
    return rnd && b.worker_types[env] == HA_ADR_WORKER_BOUNDARY;
  };
  int cnt = 0;
  for (int env = beg + tid; env < end; env += kAdrThreads) cnt += qualifies(env);
  cnt = warp_sum_i(cnt);
  if (lane == 0) warp_cnt[warp] = cnt;
  __syncthreads();
  if (tid == 0) {
    int total = 0;
#pragma unroll
    for (int w = 0; w < kAdrThreads / 32; ++w) total += warp_cnt[w];
    *reinterpret_cast<volatile long long*>(b.block_counts + blockIdx.x) = (epoch << 32) | (long long)total;
  }
  long long pre = 0, tot = 0;
  for (int i = tid; i < (int)gridDim.x; i += kAdrThreads) {
    long long v;
    do { v = adr_ld_volatile(b.block_counts + i); } while ((v >> 32) != epoch);
    const long long c = v & 0xffffffffLL;
    tot += c;
    if (i < (int)blockIdx.x) pre += c;
  }
  pre = warp_sum_ll(pre);
  tot = warp_sum_ll(tot);
  __syncthreads();
  if (lane == 0) { red[0][warp] = pre; red[1][warp] = tot; }
  __syncthreads();
  pre = 0; tot = 0;
#pragma unroll
  for (int w = 0; w < kAdrThreads / 32; ++w) { pre += red[0][w]; tot += red[1][w]; }
  if (tot > 0) {
    long long running = pre;
    for (int base = beg; base < end; base += kAdrThreads) {
      const int env = base + tid;
      const bool q = qualifies(env);
      const unsigned m = __ballot_sync(0xffffffffu, q);
      if (lane == 0) warp_cnt[warp] = __popc(m);
      __syncthreads();
      int before = 0, sub_total = 0;
#pragma unroll
      for (int w = 0; w < kAdrThreads / 32; ++w) { before += (w < warp) ? warp_cnt[w] : 0; sub_total += warp_cnt[w]; }
      if (q) b.list[running + before + __popc(m & ((1u << lane) - 1u))] = env;
      running += sub_total;
      __syncthreads();
    }
  }
  if (tid == 0) {
    __threadfence();
    const unsigned long long t = atomicAdd(reinterpret_cast<unsigned long long*>(b.state + 2), 1ULL);
    if (t == (unsigned long long)gridDim.x - 1ULL) {
      b.state[0] = tot;            // length of the list for launch 2
      b.state[1] = epoch;
      b.state[2] = 0;
    }
  }
}

// Launch 2, CTA = queue q (mode index 2n / 2n+1): walk the list in order, extend the ring with the objectives of the envs whose
// adr_mode is q (deque.extend: entries beyond maxlen push out the oldest), then mean (float64, as np.mean of the deque) and length.
__global__ void __launch_bounds__(128)
adr_extend_kernel(const int cap, const __grid_constant__ HaAdrQueueBuffers b) {
  __shared__ int warp_cnt[4];
  __shared__ double s_sum[4];
  const int q = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int total = (int)b.state[0];
  long long appended = b.appended[q];
  float* ring = b.queues + (size_t)q * cap;
  for (int base = 0; base < total; base += 128) {
    const int i = base + tid;
    int env = 0;
    bool mine = false;
    if (i < total) { env = b.list[i]; mine = b.adr_modes[env] == (long long)q; }
    const unsigned m = __ballot_sync(0xffffffffu, mine);
    if (lane == 0) warp_cnt[warp] = __popc(m);
    __syncthreads();
    int before = 0, sub_total = 0;
#pragma unroll
    for (int w = 0; w < 4; ++w) { before += (w < warp) ? warp_cnt[w] : 0; sub_total += warp_cnt[w]; }
    // entries of one 128-wide slab never collide in the ring as long as cap >= 128; otherwise only the survivors write
    if (mine) {
      const long long r = before + __popc(m & ((1u << lane) - 1u));
      if (r >= (long long)sub_total - cap) ring[(appended + r) % cap] = b.objective[env];
    }
    appended += sub_total;
    __syncthreads();
  }
  const int valid = (int)(appended < (long long)cap ? appended : (long long)cap);
  double sum = 0.0;
  for (int i = tid; i < valid; i += 128) sum += (double)ring[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) s_sum[warp] = sum;
  __syncthreads();
  if (tid == 0) {
    sum = s_sum[0] + s_sum[1] + s_sum[2] + s_sum[3];
    b.appended[q] = appended;
    b.summary[2 * q + 0] = valid ? sum / (double)valid : 0.0;      // mean_low / mean_high (0 for an empty queue, :804-805)
    b.summary[2 * q + 1] = (double)valid;                          // len(queue)
  }
}

// ---- scratch of ha_adr_queue_step_static: lives in the module image, so every GPU that loads the library holds it until exit ----
struct AdrStaticScratch {
  int64_t state[4];                              // scan state (list length, epoch, ticket, unused)
  int64_t block_counts[HA_OBSERVER_MAX_CTAS];    // per-CTA counts tagged with the epoch
  int32_t list[HA_ADR_STATIC_MAX_ENVS];          // finished boundary workers in env order
};
constexpr size_t kAdrStaticZeroBytes = offsetof(AdrStaticScratch, list);   // state + block_counts must start at zero for every call

}  // namespace ha

// File scope.  In host code this name is only a stand-in for the symbol (NOT a device address): the device address comes from
// cudaGetSymbolAddress below, once per device.
static __device__ __align__(256) unsigned char g_adr_static_scratch[sizeof(ha::AdrStaticScratch)];

namespace {

int adr_queue_launch(int32_t n_envs, int32_t n_queues, int32_t capacity, int32_t all_envs, const HaAdrQueueBuffers& b, cudaStream_t stream) {
  int grid = (n_envs + ha::kAdrThreads - 1) / ha::kAdrThreads;
  if (grid > HA_OBSERVER_MAX_CTAS) grid = HA_OBSERVER_MAX_CTAS;
  int per_cta = (n_envs + grid - 1) / grid;
  per_cta = (per_cta + ha::kAdrThreads - 1) / ha::kAdrThreads * ha::kAdrThreads;
  grid = (n_envs + per_cta - 1) / per_cta;
  int rc = ha::check_cuda(ha::launch(ha::adr_compact_kernel, grid, ha::kAdrThreads, 0, stream, (int)n_envs, per_cta, (int)all_envs, b),
                          "adr_compact_kernel launch");
  if (rc) return rc;
  return ha::check_cuda(ha::launch(ha::adr_extend_kernel, (int)n_queues, 128, 0, stream, (int)capacity, b), "adr_extend_kernel launch");
}

// device address of the static scratch on the CURRENT device; resolved once per device, with that device current
int adr_static_scratch(ha::AdrStaticScratch** out) {
  static ha::AdrStaticScratch* addr[ha::kMaxDevices] = {};
  int dev = -1;
  int rc = ha::check_cuda(cudaGetDevice(&dev), "cudaGetDevice");
  if (rc) return rc;
  HA_REQUIRE(dev >= 0 && dev < ha::kMaxDevices, "device ordinal %d beyond the %d the library keeps per-device state for", dev, ha::kMaxDevices);
  if (addr[dev] == nullptr) {
    void* ptr = nullptr;
    rc = ha::check_cuda(cudaGetSymbolAddress(&ptr, g_adr_static_scratch), "cudaGetSymbolAddress(g_adr_static_scratch)");
    if (rc) return rc;
    HA_REQUIRE(ptr != nullptr, "cudaGetSymbolAddress returned a null address for the static scratch");
    addr[dev] = static_cast<ha::AdrStaticScratch*>(ptr);
  }
  *out = addr[dev];
  return HA_OK;
}

}  // namespace

extern "C" {

size_t ha_sizeof_adr_queue_buffers(void) { return sizeof(HaAdrQueueBuffers); }

int ha_adr_queue_step_static(int32_t n_envs, int32_t n_queues, int32_t capacity, int32_t all_envs, const HaAdrQueueBuffers* b, ha_stream_t stream) {
  HA_REQUIRE(b != nullptr, "null buffers");
  HA_REQUIRE(n_envs > 0 && n_queues > 0 && capacity > 0, "n_envs, n_queues and capacity must be positive");
  HA_REQUIRE(n_envs <= HA_ADR_STATIC_MAX_ENVS,
             "ha_adr_queue_step_static: n_envs %d exceeds HA_ADR_STATIC_MAX_ENVS (%d), the size of the library's static scratch; "
             "use ha_adr_queue_step with caller-provided scratch", n_envs, HA_ADR_STATIC_MAX_ENVS);
  HA_REQUIRE(b->reset_buf && b->worker_types && b->adr_modes && b->objective && b->queues && b->appended && b->summary,
             "HaAdrQueueBuffers has a null input / output field (list, state and block_counts are ignored here)");
  ha::AdrStaticScratch* scratch = nullptr;
  int rc = adr_static_scratch(&scratch);
  if (rc) return rc;
  HaAdrQueueBuffers local = *b;                    // inputs and outputs stay the caller's; only the scratch pointers are ours
  local.state = scratch->state;
  local.block_counts = scratch->block_counts;
  local.list = scratch->list;
  // the array keeps the last call's contents: the scan state and the tagged counts must start at zero (the list is fully
  // rewritten up to the length the first launch leaves in state[0]).  Stream-ordered, no allocation.
  rc = ha::check_cuda(cudaMemsetAsync(scratch, 0, ha::kAdrStaticZeroBytes, (cudaStream_t)stream), "cudaMemsetAsync(static scratch)");
  if (rc) return rc;
  return adr_queue_launch(n_envs, n_queues, capacity, all_envs, local, (cudaStream_t)stream);
}

int ha_adr_queue_step(int32_t n_envs, int32_t n_queues, int32_t capacity, int32_t all_envs, const HaAdrQueueBuffers* b, ha_stream_t stream) {
  HA_REQUIRE(b != nullptr, "null buffers");
  HA_REQUIRE(n_envs > 0 && n_queues > 0 && capacity > 0, "n_envs, n_queues and capacity must be positive");
  HA_REQUIRE(b->reset_buf && b->worker_types && b->adr_modes && b->objective && b->queues && b->appended && b->list && b->state &&
             b->block_counts && b->summary, "HaAdrQueueBuffers has a null field");
  int grid = (n_envs + ha::kAdrThreads - 1) / ha::kAdrThreads;
  if (grid > HA_OBSERVER_MAX_CTAS) grid = HA_OBSERVER_MAX_CTAS;
  int per_cta = (n_envs + grid - 1) / grid;
  per_cta = (per_cta + ha::kAdrThreads - 1) / ha::kAdrThreads * ha::kAdrThreads;
  grid = (n_envs + per_cta - 1) / per_cta;
  int rc = ha::check_cuda(ha::launch(ha::adr_compact_kernel, grid, ha::kAdrThreads, 0, (cudaStream_t)stream, (int)n_envs, per_cta, (int)all_envs, *b),
                          "adr_compact_kernel launch");
  if (rc) return rc;
  return ha::check_cuda(ha::launch(ha::adr_extend_kernel, (int)n_queues, 128, 0, (cudaStream_t)stream, (int)capacity, *b), "adr_extend_kernel launch");
}

}  // extern "C"
