// Device-side episode_cumulative bookkeeping of the rl_games observer (rlgames_utils.py:149-172, :192-205); see
// include/handarm_b200.h.  One streaming launch per step from the done mask: no nonzero(), no per-env .item().
#include <cuda_runtime.h>
#include <math.h>

#include "../../include/handarm_b200.h"
#include "ha_device.cuh"
#include "ha_host.h"

namespace ha {

constexpr int kObsThreads = 256;

__device__ __forceinline__ long long ld_volatile(const int64_t* p) { return *reinterpret_cast<const volatile long long*>(p); }

// CTA = a contiguous chunk of envs (thread = env, kObsThreads at a time).
//   pass 1  cumulative[k][env] += terms[k][env] for every env (coalesced per term row); count the chunk's finished envs
//   scan    every CTA publishes its count tagged with the launch epoch and reads the counts of all CTAs (the grid is small and
//           co-resident): exclusive prefix = deque position of its first finished env, total = episodes finished this step
//   pass 2  finished env with rank r (ascending env order, as done_indices lists them): if the deque still holds it after
//           this step's appends (r >= total - capacity), ring[k][(appended + r) % capacity] = cumulative[k][env]; then the
//           running sums of the env restart at 0
// The last CTA (ticket) advances the append counter and the epoch.
__global__ void __launch_bounds__(kObsThreads)
observer_step_kernel(const int n, const int nt, const int cap, const int per_cta, const __grid_constant__ HaObserverBuffers b) {
  __shared__ int warp_cnt[kObsThreads / 32];
  __shared__ long long red[2][kObsThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long head = ld_volatile(b.state + 0);
  const long long epoch = ld_volatile(b.state + 1) + 1;
  const int beg = blockIdx.x * per_cta, end = min(n, beg + per_cta);

  int cnt = 0;
  for (int env = beg + tid; env < end; env += kObsThreads) {
    cnt += (b.dones[env] != 0);
    for (int k = 0; k < nt; ++k) {
      const size_t i = (size_t)k * n + env;
      b.cumulative[i] = b.cumulative[i] + __ldg(b.terms + i);     // self.episode_cumulative[key] += value
    }
  }
  cnt = warp_sum_i(cnt);
  if (lane == 0) warp_cnt[warp] = cnt;
  __syncthreads();
  if (tid == 0) {
    int total = 0;
#pragma unroll
    for (int w = 0; w < kObsThreads / 32; ++w) total += warp_cnt[w];
    *reinterpret_cast<volatile long long*>(b.block_counts + blockIdx.x) = (epoch << 32) | (long long)total;
  }
  // counts of all CTAs of this launch
  long long pre = 0, tot = 0;
  for (int i = tid; i < (int)gridDim.x; i += kObsThreads) {
    long long v;
    do { v = ld_volatile(b.block_counts + i); } while ((v >> 32) != epoch);
    const long long c = v & 0xffffffffLL;
    tot += c;
    if (i < (int)blockIdx.x) pre += c;
  }
  pre = warp_sum_ll(pre);
  tot = warp_sum_ll(tot);
  __syncthreads();                                      // warp_cnt is reused below
  if (lane == 0) { red[0][warp] = pre; red[1][warp] = tot; }
  __syncthreads();
  pre = 0; tot = 0;
#pragma unroll
  for (int w = 0; w < kObsThreads / 32; ++w) { pre += red[0][w]; tot += red[1][w]; }

  if (tot > 0) {                                        // uniform over the grid
    long long running = pre;
    for (int base = beg; base < end; base += kObsThreads) {
      const int env = base + tid;
      const bool done = env < end && b.dones[env] != 0;
      const unsigned m = __ballot_sync(0xffffffffu, done);
      if (lane == 0) warp_cnt[warp] = __popc(m);
      __syncthreads();
      int before = 0, sub_total = 0;
#pragma unroll
      for (int w = 0; w < kObsThreads / 32; ++w) { before += (w < warp) ? warp_cnt[w] : 0; sub_total += warp_cnt[w]; }
      if (done) {
        const long long r = running + before + __popc(m & ((1u << lane) - 1u));
        const bool kept = r >= tot - (long long)cap;     // deque(maxlen): later appends of this step push it out otherwise
        const long long slot = (head + r) % (long long)cap;
        for (int k = 0; k < nt; ++k) {
          const size_t i = (size_t)k * n + env;
          if (kept) b.ring[(size_t)k * cap + slot] = b.cumulative[i];   // episode_cumulative_avg[key].append(cum[key][done_idx])
          b.cumulative[i] = 0.0f;                                       // self.episode_cumulative[key][done_idx] = 0
        }
      }
      running += sub_total;
      __syncthreads();
    }
  }
  // last CTA: the deque grew by `tot`, next launch gets a new epoch
  if (tid == 0) {
    __threadfence();
    const unsigned long long t = atomicAdd(reinterpret_cast<unsigned long long*>(b.state + 2), 1ULL);
    if (t == (unsigned long long)gridDim.x - 1ULL) {
      b.state[0] = head + tot;
      b.state[1] = epoch;
      b.state[2] = 0;
    }
  }
}

// CTA = one term: mean / min / max over the valid part of its ring (np.mean / np.min / np.max of the deque, :199-201)
__global__ void __launch_bounds__(128) observer_summary_kernel(const int nt, const int cap, const __grid_constant__ HaObserverBuffers b) {
  __shared__ double s_sum[4];
  __shared__ float s_mn[4], s_mx[4];
  const int k = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long long appended = b.state[0];
  const int valid = (int)(appended < (long long)cap ? appended : (long long)cap);
  double sum = 0.0;
  float mn = INFINITY, mx = -INFINITY;
  for (int i = tid; i < valid; i += 128) {
    const float v = b.ring[(size_t)k * cap + i];
    sum += (double)v;
    mn = fminf(mn, v);
    mx = fmaxf(mx, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  }
  if (lane == 0) { s_sum[warp] = sum; s_mn[warp] = mn; s_mx[warp] = mx; }
  __syncthreads();
  if (tid == 0) {
    for (int w = 1; w < 4; ++w) { sum += s_sum[w]; mn = fminf(mn, s_mn[w]); mx = fmaxf(mx, s_mx[w]); }
    const float nanv = __int_as_float(0x7fc00000);
    b.summary[0 * nt + k] = valid ? (float)(sum / (double)valid) : nanv;
    b.summary[1 * nt + k] = valid ? mn : nanv;
    b.summary[2 * nt + k] = valid ? mx : nanv;
  }
}

}  // namespace ha

extern "C" {

size_t ha_sizeof_observer_buffers(void) { return sizeof(HaObserverBuffers); }

int ha_observer_step(int32_t n_envs, int32_t n_terms, int32_t capacity, const HaObserverBuffers* b, ha_stream_t stream) {
  HA_REQUIRE(b != nullptr, "null buffers");
  HA_REQUIRE(n_envs > 0, "n_envs must be positive (got %d)", n_envs);
  HA_REQUIRE(n_terms > 0 && n_terms <= HA_OBSERVER_MAX_TERMS, "n_terms must be in 1..%d (got %d)", HA_OBSERVER_MAX_TERMS, n_terms);
  HA_REQUIRE(capacity > 0, "capacity (games_to_track) must be positive (got %d)", capacity);
  HA_REQUIRE(b->terms && b->dones && b->cumulative && b->ring && b->state && b->block_counts, "HaObserverBuffers has a null field");
  int grid = (n_envs + ha::kObsThreads - 1) / ha::kObsThreads;
  if (grid > HA_OBSERVER_MAX_CTAS) grid = HA_OBSERVER_MAX_CTAS;      // the scan spins on every CTA's count: the grid stays co-resident
  int per_cta = (n_envs + grid - 1) / grid;
  per_cta = (per_cta + ha::kObsThreads - 1) / ha::kObsThreads * ha::kObsThreads;
  grid = (n_envs + per_cta - 1) / per_cta;
  return ha::check_cuda(ha::launch(ha::observer_step_kernel, grid, ha::kObsThreads, 0, (cudaStream_t)stream, (int)n_envs, (int)n_terms,
                                   (int)capacity, per_cta, *b), "observer_step_kernel launch");
}

int ha_observer_summary(int32_t n_terms, int32_t capacity, const HaObserverBuffers* b, ha_stream_t stream) {
  HA_REQUIRE(b != nullptr, "null buffers");
  HA_REQUIRE(n_terms > 0 && n_terms <= HA_OBSERVER_MAX_TERMS, "n_terms must be in 1..%d (got %d)", HA_OBSERVER_MAX_TERMS, n_terms);
  HA_REQUIRE(capacity > 0, "capacity (games_to_track) must be positive (got %d)", capacity);
  HA_REQUIRE(b->ring && b->state && b->summary, "HaObserverBuffers has a null field");
  return ha::check_cuda(ha::launch(ha::observer_summary_kernel, (int)n_terms, 128, 0, (cudaStream_t)stream, (int)n_terms, (int)capacity, *b),
                        "observer_summary_kernel launch");
}

}  // extern "C"
