// Cross-rank episode statistics over NVLink peer memory, device side; see include/handarm_b200.h (HaStatsExchange).
// One warp runs a whole exchange: push this rank's 16 floats + a sequence tag into every peer's buffer with plain remote
// stores, and reduce the newest exchange whose vectors have all arrived.  Wait-free on both sides.  The same function is the
// body of the single-warp launch ha_stats_exchange and of the in-kernel exchange of the post-physics kernel (no launch at all).
#pragma once
#include <cuda_runtime.h>

#include "../../include/handarm_b200.h"

namespace ha {

// exchange buffer of one rank: unsigned long long words[RING][world][HA_STATS_LEN], long long head[world].
// A word carries one statistic AND the number of the exchange it belongs to: (exchange number << 32) | float bits.  One aligned
// 8-byte store is atomic, so a reader can never see a value without its tag: no fences, no release / acquire, no second look.
__host__ __device__ inline size_t xchg_words(int world) { return (size_t)HA_XCHG_RING * world * HA_STATS_LEN; }

__device__ __forceinline__ void st_relaxed_sys_u64(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys_u64(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}

// Protocol (exchange number seq, slot seq % RING of every peer).
//   Reader first: heads of all ranks -> m = min = the newest exchange every rank CLAIMS to have published (a hint: heads and
//   words travel unordered); if it is newer than the last one reduced and its slot is not about to be overwritten, read the W
//   words of every statistic; the exchange counts as complete only if every word carries number m (a word still in flight, or
//   already overwritten by a newer exchange, carries another number).  Reduced in rank order (the same bits on every rank);
//   otherwise `reduced` keeps its value and the next call looks again.
//   Resync: a rank that finds itself a whole ring behind the fastest one can never see a complete slot again (its peers have
//   overwritten them), so it jumps its exchange number to the largest published one (status |= 1) instead of counting on.
//   Writer: 16 tagged words into slot seq % RING of every peer, head[rank] = seq on every peer.  Plain remote stores, nothing to
//   wait for.
// Nobody ever waits: a rank that is late only delays the exchange number the others reduce; while a peer is a whole ring behind
// the others skip and raise status |= 2.
__device__ __forceinline__ void stats_exchange_warp(const HaStatsExchange& x, const float* __restrict__ stats, float* __restrict__ reduced,
                                                    const int lane) {
  const unsigned FULL = 0xffffffffu;
  const int W = x.world;
  const float mine_v = lane < HA_STATS_LEN ? stats[lane] : 0.f;
  const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(x.peer_bufs[x.rank]);
  const long long* head = reinterpret_cast<const long long*>(mine + xchg_words(W));
  long long seq = x.seq[0] + 1;
  const long long last = x.seq[1];
  // ---- heads: newest complete exchange (min) and the fastest rank (max) ----
  long long h = 0x7fffffffffffffffLL, hx = 0;
  if (lane < W) { h = (long long)ld_relaxed_sys_u64(reinterpret_cast<const unsigned long long*>(head + lane)); hx = h; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const long long t = __shfl_xor_sync(FULL, h, o), u = __shfl_xor_sync(FULL, hx, o);
    h = t < h ? t : h;
    hx = u > hx ? u : hx;
  }
  int status_bits = 0;
  if (hx - seq >= HA_XCHG_RING - 1) { seq = hx; status_bits |= 1; }      // lapped: rejoin at the fastest rank's number
  const long long m = h;
  const bool newer = m >= 1 && m > last;
  bool ok = newer && seq - m < HA_XCHG_RING - 1;
  if (newer && !ok && !(status_bits & 1)) status_bits |= 2;               // a peer is a whole ring behind
  const int pr = (int)((m > 0 ? m : 0) % HA_XCHG_RING);
  float acc = 0.f;
  bool all_tagged = true;
  if (ok && lane < HA_STATS_LEN) {
    const bool is_max = (x.max_mask >> lane) & 1u;
    const unsigned long long* slot = mine + (size_t)pr * W * HA_STATS_LEN + lane;
    for (int r = 0; r < W; ++r) {                        // rank order: the same association on every rank
      const unsigned long long w = ld_relaxed_sys_u64(slot + (size_t)r * HA_STATS_LEN);
      all_tagged = all_tagged && (unsigned)(w >> 32) == (unsigned)m;
      const float v = __uint_as_float((unsigned)(w & 0xffffffffULL));
      acc = r == 0 ? v : (is_max ? fmaxf(acc, v) : acc + v);
    }
  }
  ok = ok && __all_sync(FULL, all_tagged);
  if (ok && lane < HA_STATS_LEN) reduced[lane] = acc;
  // ---- writer ----
  const int ring = (int)(seq % HA_XCHG_RING);
  if (lane < HA_STATS_LEN) {
    const unsigned long long word = ((unsigned long long)(unsigned)seq << 32) | (unsigned long long)__float_as_uint(mine_v);
    for (int peer = 0; peer < W; ++peer)
      st_relaxed_sys_u64(reinterpret_cast<unsigned long long*>(x.peer_bufs[peer]) + ((size_t)ring * W + x.rank) * HA_STATS_LEN + lane, word);
  }
  if (lane < W) {
    unsigned long long* peer_head = reinterpret_cast<unsigned long long*>(x.peer_bufs[lane]) + xchg_words(W);
    st_relaxed_sys_u64(peer_head + x.rank, (unsigned long long)seq);
  }
  if (lane == 0) {
    x.seq[0] = seq;
    if (ok) x.seq[1] = m;
    if (status_bits) *x.status = *x.status | status_bits;
  }
}

}  // namespace ha
