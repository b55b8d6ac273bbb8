// K_post for the single-arm AllegroKuka family: one launch replaces
// allegro_kuka_base.py:1426-1443 (post_physics_step) = compute_observations
// (:991-1089) + compute_full_state (:1091-1172) + compute_kuka_reward (:854-930)
// + _compute_resets (:841-849) + clamp_obs (:1174-1176), plus the VecTask.step
// tail (vec_task.py:424-432: timeout_buf, obs clip).
//
// Mapping (memory-bound: ~1.4 KB of 128-byte lines fetched and 613 B written per env):
//   CTA = 4 warps working on one tile of 32 consecutive envs at a time; CTAs are
//   persistent (grid = resident CTAs) and walk tiles blockIdx.x, +gridDim.x, ...
//   load     every input that is not a per-env scalar lands in shared memory through
//            cp.async (LDGSTS): 4-byte copies for the AoS rows (13-float rigid-body /
//            root rows sit at 52-byte stride, only 4-byte aligned), 16-byte copies for
//            the contiguous tiles (DOF state, keypoint table, scales).  Two stages: the
//            next tile's copies are in flight while the current tile is computed and
//            stored, so DRAM reads, the arithmetic and the stores overlap.
//   compute  lane = env in every warp, the warps split the work by role:
//              warp 0  palm centre + 4 fingertips            -> tip_rel_palm, distances
//              warp 1  K object / goal keypoints              -> kp_rel_palm, kp_rel_goal, kp_max
//              warp 2  DOF tile: unscale(pos), vel, |vel| sums
//              warp 3  owns the per-env state (registers): row copies, then after a CTA
//                      barrier the reward / reset logic, per-env stores, statistics
//   store    the 32 x obs_dim tile leaves shared memory as one contiguous run of
//            128-bit stores issued by all four warps.
// Episode statistics: per-lane 64-bit fixed-point partials, reduced once per CTA with
// integer shuffles, so every sum is order-independent and bit-reproducible for any
// grid size.  Single arm: each CTA leaves its ten partials in its own slot of
// stat_accum with plain stores (no atomics, no ticket, nothing the kernel's tail waits
// for: the atomics + ticket + last-CTA epilogue cost 2.8 us of a 13.1 us launch on B200);
// the slots are folded into stats[] by the next launch of the step loop (K_pre, CTA 0)
// or on demand by ha_kuka_stats_finalize.  Two arms (non-persistent grid): integer
// atomics + last-CTA ticket inside the launch (kuka_post_stats).
#pragma once
#include "ha_device.cuh"
#include "../../include/handarm_b200.h"
#include "ha_stats_xchg.cuh"

// tuning switches (A/B builds): which transfers of K_post go through the bulk-copy engine (cp.async.bulk, SASS UBLKCP).
// Measured on B200, 32768 envs, component planes (K_post us): none 13.4 | store 13.9 | images + store 13.8 | rows + images + store 15.0.
// Also measured: hoisting the plane-row decoding of the cp.async issue into a per-thread copy plan (12 more live registers) 13.0 -> 14.1.
// 128-byte rows are too small for the copy engine (57 descriptors per tile against 8 LDGSTS per thread), and a bulk store needs the
// +-clamp moved from the store loop into the role warps, which lengthens the per-tile critical path by more than the store saves.
#ifndef HA_BULK_ROWS
#define HA_BULK_ROWS 0      // the 54 component-plane rows of a tile (128 bytes each)
#endif
#ifndef HA_BULK_IMAGES
#define HA_BULK_IMAGES 0    // the contiguous keypoint / scale / DOF images
#endif
#ifndef HA_BULK_STORE
#define HA_BULK_STORE 0     // the finished [32, OBS] observation tile
#endif
#ifndef HA_LATE_PREFETCH
#define HA_LATE_PREFETCH 0  // 1: issue the next tile's copies only after this tile's inputs have landed (measured: 11.3 vs 11.0 us, worse)
#endif
#ifndef HA_EARLY_STORE
#define HA_EARLY_STORE 1    // observation words without late columns leave before the reward chain has finished
#endif
#ifndef HA_POST_NS1_CTAS
#define HA_POST_NS1_CTAS 7   // resident single-stage CTAs per SM the register allocation is sized for
#endif

namespace ha {

struct KukaLayout {            // obs slot offsets, allegro_kuka_base.py:1091-1172
  int dof_pos, dof_vel, palm_pos, palm_rot_vel, obj_rot_vel, tip_rel_palm, kp_rel_palm, kp_rel_goal,
      scales, closest_kp, closest_tip, lifted, progress, successes, reward, dim;
};
__host__ __device__ constexpr KukaLayout kuka_layout(int n_dofs, int n_tips, int K) {
  KukaLayout l{};
  int o = 0;
  l.dof_pos = o; o += n_dofs;
  l.dof_vel = o; o += n_dofs;
  l.palm_pos = o; o += 3;
  l.palm_rot_vel = o; o += 10;
  l.obj_rot_vel = o; o += 10;
  l.tip_rel_palm = o; o += 3 * n_tips;
  l.kp_rel_palm = o; o += 3 * K;
  l.kp_rel_goal = o; o += 3 * K;
  l.scales = o; o += 3;
  l.closest_kp = o; o += 1;
  l.closest_tip = o; o += n_tips;
  l.lifted = o; o += 1;
  l.progress = o; o += 1;
  l.successes = o; o += 1;
  l.reward = o; o += 1;
  l.dim = o;
  return l;
}

enum { kPostFull = 0, kPostObserveOnly = 1 };

constexpr int kD1 = 23;              // single arm DOFs
constexpr int kT1 = 4;               // fingertips
constexpr int kTile = 32;            // envs per tile (lane = env)
constexpr int kPostThreads = 128;    // 4 role warps
constexpr int kNPart = HA_KUKA_N_PARTIAL_STATS;

__host__ __device__ constexpr bool stat_is_max(int i) {
  return i == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE || i == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE;
}
// fixed-point scale of the 64-bit statistic accumulators
__host__ __device__ constexpr float stat_scale(int i) {
  return (i == HA_KUKA_STAT_SUM_REWARD) ? 65536.0f
         : (i == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE || i == HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES) ? 1048576.0f : 1.0f;
}
constexpr long long kStatMaxInit = -0x7fffffffLL - 1;   // below every float key

// One tile's inputs as they land in shared memory.  Row strides are odd (13, 29, 7) so lane = env reads are
// conflict-free; the contiguous images (kp, dof) are read with 128 / 64-bit loads, which are conflict-free at
// their 12- and 46-word strides.
template <int K>
struct KukaPostStage {
  static constexpr int kTipW = kT1 * 7 + 1;       // 29
  float palm[kTile * 13];                          // [e][13]  pos3 quat4 linvel3 angvel3
  float obj[kTile * 13];                           // [e][13]
  float tip[kTile * kTipW];                        // [e][t * 7 + c]  pos3 quat4
  float goal[kTile * 13];                          // [e][13] contiguous image of goal_states rows (pos3 quat4 used)
  float kp[kTile * 3 * K];                         // contiguous image of object_keypoint_offsets[env0 .. env0+31]
  float scales[kTile * 3];                         // contiguous image
  float dof[kTile * 2 * kD1];                      // contiguous image of dof_state (pos, vel pairs)
};

template <int K, int NS>
struct KukaPostSmem {
  static constexpr int kObs = kuka_layout(kD1, kT1, K).dim;
  KukaPostStage<K> st[NS];
  float obs[kTile * kObs];            // row-major tile, leaves as one contiguous run
  float xch[8 * kTile];               // [q][e]: cur[0..3], kp_max, sum|vel| arm, sum|vel| hand
  float terms[HA_KUKA_N_EPISODE_TERMS * kTile];   // [k][e]: this step's reward terms, state warp -> the warps that keep the per-episode sums
  unsigned long long mbar[2];         // transaction barriers of the bulk-copied stages
  const float* rowp[13 + 13 + kT1 * 7];   // component planes: source of staged row r at env 0 (palm 0..12, object 13..25, tips 26..53)
};
// a tile whose inputs arrive by bulk copies: component planes, full tile, 16-byte aligned plane rows
__device__ __forceinline__ bool kuka_post_bulk_tile(bool soa, int n_envs, int env0) {
  return (HA_BULK_ROWS || HA_BULK_IMAGES) && soa && (n_envs & 3) == 0 && n_envs - env0 >= kTile;
}
static_assert(sizeof(KukaPostStage<4>) % 16 == 0 && sizeof(KukaPostStage<1>) % 16 == 0, "stage must keep 16-byte alignment");
static_assert((kTile * 13 * 2 + kTile * 29 + kTile * 13) % 4 == 0, "kp image must be 16-byte aligned inside the stage");

// per-lane statistic partials carried across the tiles of a persistent CTA (warp 3 only)
struct KukaPostAcc {
  long long rew = 0, prev_ep = 0, tobj = 0;       // fixed point
  float mx = -3.0e38f, neg_mn = -3.0e38f;
  int n = 0, resets = 0, success = 0, timeout = 0, lifted = 0;
};

// Episode statistics of one CTA -> global fixed-point accumulators; the last CTA of the grid converts them
// to floats.  Called once per CTA by warp 3 after its last tile.
__device__ __forceinline__ void kuka_post_stats(const HaKukaBuffers& b, const KukaPostAcc& a, const int lane) {
  const unsigned FULL = 0xffffffffu;
  const long long s_rew = warp_sum_ll(a.rew), s_pe = warp_sum_ll(a.prev_ep), s_to = warp_sum_ll(a.tobj);
  const int s_n = warp_sum_i(a.n), s_rs = warp_sum_i(a.resets), s_su = warp_sum_i(a.success), s_tm = warp_sum_i(a.timeout),
            s_li = warp_sum_i(a.lifted);
  const float v_mx = warp_max(a.mx), v_mn = warp_max(a.neg_mn);
  long long* accum = reinterpret_cast<long long*>(b.stat_accum);
  int is_last = 0;
  {
    long long val = 0;
    // selects, not a switch: a switch on the lane index is a divergent jump table (one serial pass per case)
    val = lane == HA_KUKA_STAT_N ? (long long)(s_n) : val;
    val = lane == HA_KUKA_STAT_SUM_REWARD ? (long long)(s_rew) : val;
    val = lane == HA_KUKA_STAT_SUM_RESETS ? (long long)(s_rs) : val;
    val = lane == HA_KUKA_STAT_SUM_SUCCESS_EVENTS ? (long long)(s_su) : val;
    val = lane == HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES ? (long long)(s_pe) : val;
    val = lane == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE ? (long long)(s_to) : val;
    val = lane == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE ? (long long)(float_key(v_mx)) : val;
    val = lane == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE ? (long long)(float_key(v_mn)) : val;
    val = lane == HA_KUKA_STAT_SUM_TIMEOUTS ? (long long)(s_tm) : val;
    val = lane == HA_KUKA_STAT_SUM_LIFTED ? (long long)(s_li) : val;
    // The accumulator atomics return their old values; consuming those values orders them (performed at L2) ahead of
    // the ticket below without a membar that would also wait for this warp's ~35 per-env stores.
    long long old = 0;
    if (lane < kNPart) {
      if (stat_is_max(lane)) old = atomicMax(accum + lane, val);
      else old = (long long)atomicAdd(reinterpret_cast<unsigned long long*>(accum + lane), (unsigned long long)val);
    }
    const unsigned dep = __any_sync(FULL, old == 0x7fffffffffffffffLL) ? 1u : 0u;   // never true; makes the ticket depend on every return
    if (lane == 0) is_last = (atomicAdd(b.tickets + 0, 1u + dep) == gridDim.x - 1);
  }
  is_last = __shfl_sync(FULL, is_last, 0);
  if (!is_last) return;
  // ---------------------- last CTA: final statistics ----------------------
  __threadfence();
  float val = 0.f;
  if (lane < kNPart) {
    const bool is_max = stat_is_max(lane);
    const long long raw = (long long)atomicExch(reinterpret_cast<unsigned long long*>(accum + lane),
                                                (unsigned long long)(is_max ? kStatMaxInit : 0LL));   // read and re-arm
    val = is_max ? key_float(raw) : (float)((double)raw / (double)stat_scale(lane));
    b.stats[lane] = val;
  }
  // means / min for extras (allegro_kuka_base.py:903-910), so the host needs no extra launches
  const float cnt = __shfl_sync(FULL, val, HA_KUKA_STAT_N);
  const float s_rew_f = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_REWARD), s_pe_f = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES);
  const float s_to_f = __shfl_sync(FULL, val, HA_KUKA_STAT_SUM_TRUE_OBJECTIVE), s_mn_f = __shfl_sync(FULL, val, HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE);
  if (lane == 0) {
    const float inv_n = 1.0f / cnt;
    b.stats[HA_KUKA_STAT_MEAN_REWARD] = s_rew_f * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES] = s_pe_f * inv_n;
    b.stats[HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE] = s_to_f * inv_n;
    b.stats[HA_KUKA_STAT_MIN_TRUE_OBJECTIVE] = -s_mn_f;
    b.tickets[0] = 0;                    // re-arm for the next launch (graph replay safe)
    // the reset/index counters of the pre-step are consumed by now: clear for the next pre-step
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    // fast-mode RNG: one counter tick per step
    b.dynamic->rng_step = b.dynamic->rng_step + 1;
  }
}

// ---- deferred statistics (single arm) ------------------------------------------------------------------------------------
// stat_accum is [HA_KUKA_N_PARTIAL_STATS][HA_KUKA_STAT_SLOTS] int64: slot = CTA of the last full post step; the slots beyond the
// post kernel's grid keep the identity the host wrote once (ha_kuka_stat_accum_init).
__device__ __forceinline__ void kuka_post_store_partials(const HaKukaBuffers& b, const KukaPostAcc& a, const int lane) {
  const long long s_rew = warp_sum_ll(a.rew), s_pe = warp_sum_ll(a.prev_ep), s_to = warp_sum_ll(a.tobj);
  const int s_n = warp_sum_i(a.n), s_rs = warp_sum_i(a.resets), s_su = warp_sum_i(a.success), s_tm = warp_sum_i(a.timeout),
            s_li = warp_sum_i(a.lifted);
  const float v_mx = warp_max(a.mx), v_mn = warp_max(a.neg_mn);
  // lane s stores statistic s: selects, not a switch -- a switch on the lane index is a ten-way divergent jump table, 0.8 us at the
  // very end of every persistent CTA
  long long val = 0;
  val = lane == HA_KUKA_STAT_N ? (long long)s_n : val;
  val = lane == HA_KUKA_STAT_SUM_REWARD ? s_rew : val;
  val = lane == HA_KUKA_STAT_SUM_RESETS ? (long long)s_rs : val;
  val = lane == HA_KUKA_STAT_SUM_SUCCESS_EVENTS ? (long long)s_su : val;
  val = lane == HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES ? s_pe : val;
  val = lane == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE ? s_to : val;
  val = lane == HA_KUKA_STAT_MAX_TRUE_OBJECTIVE ? float_key(v_mx) : val;
  val = lane == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE ? float_key(v_mn) : val;
  val = lane == HA_KUKA_STAT_SUM_TIMEOUTS ? (long long)s_tm : val;
  val = lane == HA_KUKA_STAT_SUM_LIFTED ? (long long)s_li : val;
  if (lane < kNPart) reinterpret_cast<long long*>(b.stat_accum)[(size_t)lane * HA_KUKA_STAT_SLOTS + blockIdx.x] = val;
}

// Folding the per-CTA slots into stats[]: statistic s is folded by CTA s of the folding launch (K_pre's first kNPart CTAs, or the
// kNPart CTAs of kuka_stats_finalize_kernel), kThreads threads each.  Every thread has its kFoldPerThread slot loads in flight at
// once (one L2 round trip), slots no post CTA wrote hold the identity (host-initialised), so the fold is idempotent and needs no
// slot count.  Three phases so that K_pre can overlap the loads with its own tile loads and reuse its CTA barrier.
constexpr int kFoldPerThread = HA_KUKA_STAT_SLOTS / kThreads;
static_assert(HA_KUKA_STAT_SLOTS % kThreads == 0, "slot count must be a multiple of the folding CTA");
struct KukaStatFold { long long v[kFoldPerThread]; };

__device__ __forceinline__ void kuka_fold_load(const HaKukaBuffers& b, const int stat, const int tid, KukaStatFold& f) {
  const long long* part = reinterpret_cast<const long long*>(b.stat_accum) + (size_t)stat * HA_KUKA_STAT_SLOTS;
#pragma unroll
  for (int i = 0; i < kFoldPerThread; ++i) f.v[i] = part[tid + kThreads * i];
}
// warp-level fold; lane 0 of every warp leaves its value in red[warp] (shared memory, kThreads / 32 entries)
__device__ __forceinline__ void kuka_fold_warp(const KukaStatFold& f, const int stat, const int tid, long long* red) {
  const bool is_max = stat_is_max(stat);
  long long v = f.v[0];
#pragma unroll
  for (int i = 1; i < kFoldPerThread; ++i) v = is_max ? (f.v[i] > v ? f.v[i] : v) : v + f.v[i];
#pragma unroll
  for (int ofs = 16; ofs > 0; ofs >>= 1) {
    const long long o = __shfl_xor_sync(0xffffffffu, v, ofs);
    v = is_max ? (o > v ? o : v) : v + o;
  }
  if ((tid & 31) == 0) red[tid >> 5] = v;
}
// one thread writes stats[stat] of the folded value and the value derived from it for extras (allegro_kuka_base.py:903-910)
__device__ __forceinline__ void kuka_fold_store(const HaKukaBuffers& b, const int stat, const int n_envs, const long long v) {
  const bool is_max = stat_is_max(stat);
  if (is_max) {
    if (v == kStatMaxInit) return;                      // no post step has run yet
    const float x = key_float(v);
    b.stats[stat] = x;
    if (stat == HA_KUKA_STAT_NEG_MIN_TRUE_OBJECTIVE) b.stats[HA_KUKA_STAT_MIN_TRUE_OBJECTIVE] = -x;
    return;
  }
  const float x = (float)((double)v / (double)stat_scale(stat));
  b.stats[stat] = x;
  const float inv_n = 1.0f / (float)n_envs;             // HA_KUKA_STAT_N counts every env of the block
  if (stat == HA_KUKA_STAT_SUM_REWARD) b.stats[HA_KUKA_STAT_MEAN_REWARD] = x * inv_n;
  if (stat == HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES) b.stats[HA_KUKA_STAT_MEAN_PREV_EP_SUCCESSES] = x * inv_n;
  if (stat == HA_KUKA_STAT_SUM_TRUE_OBJECTIVE) b.stats[HA_KUKA_STAT_MEAN_TRUE_OBJECTIVE] = x * inv_n;
}
// after a CTA barrier: one thread combines the warps' values (kuka_fold_warp) and stores the statistic
__device__ __forceinline__ void kuka_fold_finish(const HaKukaBuffers& b, const int stat, const int n_envs, const long long* red) {
  const bool is_max = stat_is_max(stat);
  long long v = red[0];
#pragma unroll
  for (int w = 1; w < kThreads / 32; ++w) v = is_max ? (red[w] > v ? red[w] : v) : v + red[w];
  kuka_fold_store(b, stat, n_envs, v);
}

__global__ void __launch_bounds__(kThreads) kuka_stats_finalize_kernel(const __grid_constant__ HaKukaBuffers b, const int n_envs) {
  __shared__ long long red[kThreads / 32];
  KukaStatFold f;
  kuka_fold_load(b, blockIdx.x, threadIdx.x, f);
  kuka_fold_warp(f, blockIdx.x, threadIdx.x, red);
  __syncthreads();
  if (threadIdx.x == 0) kuka_fold_finish(b, blockIdx.x, n_envs, red);
}

// Issue the cp.async copies of one tile (all 128 threads).  Idle envs of a ragged last tile re-read the last valid env.
template <int K, bool SOA, int NT = kPostThreads>
__device__ __forceinline__ void kuka_post_issue(const HaKukaParams& p, const HaKukaBuffers& b, KukaPostStage<K>& s, const int env0,
                                                const int nvalid, const int tid, unsigned long long* bar, const float* const* rowp) {
  using S = KukaPostStage<K>;
  const int last = nvalid - 1;
  const bool bulk = kuka_post_bulk_tile(SOA, p.n_envs, env0);
  const size_t rbs = (size_t)p.n_bodies * 13;
  const float* rb0 = b.rigid_body_state + (size_t)env0 * rbs;
  if constexpr (SOA) {
    // component planes [body][13][n_envs]: every (body, component) row of the tile is 32 consecutive floats.
    // rows 0..12 palm, 13..25 object root, 26..53 fingertip (t, c); staged component-major [row][32].
    constexpr int kRows = 13 + 13 + kT1 * 7;
    const size_t n = (size_t)p.n_envs;
    auto row_src = [&](int r) -> const float* {
      if (r < 13) return b.rigid_body_state + ((size_t)p.palm_body[0] * 13 + r) * n + env0;
      if (r < 26) return b.root_state + ((size_t)p.object_slot * 13 + (r - 13)) * n + env0;
      const int rr = r - 26, t = rr / 7, c = rr - 7 * t;
      return b.rigid_body_state + ((size_t)p.tip_body[t] * 13 + c) * n + env0;
    };
    auto row_dst = [&](int r) -> float* { return r < 13 ? s.palm + r * kTile : r < 26 ? s.obj + (r - 13) * kTile : s.tip + (r - 26) * kTile; };
    if (bulk) {
      constexpr unsigned kBytes = (HA_BULK_ROWS ? kRows * kTile * 4 : 0) + (HA_BULK_IMAGES ? kTile * 3 * K * 4 + kTile * 3 * 4 + kTile * 2 * kD1 * 4 : 0);
      if (tid == 0) mbar_arrive_expect_tx(bar, kBytes);
    }
    if (bulk && HA_BULK_ROWS) {                         // one 128-byte bulk copy per plane row, issued by the first warp
      if (tid < 32) {
        bulk_g2s(row_dst(tid), row_src(tid), kTile * 4, bar);
        if (tid + 32 < kRows) bulk_g2s(row_dst(tid + 32), row_src(tid + 32), kTile * 4, bar);
      }
    } else if (nvalid == kTile && (p.n_envs & 3) == 0) {       // plane rows are 16-byte aligned: 8 x 16-byte copies per row
      // staged rows are contiguous (palm, obj, tip images follow each other: row r at 32 * r floats), sources come from the CTA's table
      static_assert(offsetof(S, obj) == 13 * kTile * 4 && offsetof(S, tip) == 26 * kTile * 4, "row images must be contiguous");
#pragma unroll
      for (int i = 0; i < (kRows * 8 + NT - 1) / NT; ++i) {
        const int q = tid + NT * i;
        if (q < kRows * 8) cp_async16(s.palm + 4 * q, rowp[q >> 3] + env0 + 4 * (q & 7));
      }
    } else {
      for (int q = tid; q < kRows * kTile; q += NT) {
        const int r = q >> 5, e = q & 31;
        cp_async4(row_dst(r) + e, row_src(r) + min(e, last));
      }
    }
  } else {
  {   // palm rows, 13 floats each
    const float* src = rb0 + (size_t)p.palm_body[0] * 13;
#pragma unroll
    for (int i = 0; i < (kTile * 13 + NT - 1) / NT; ++i) {
      const int j = tid + NT * i;
      if (j < kTile * 13) {
        const int e = j / 13, c = j - 13 * e;
        cp_async4(s.palm + j, src + (size_t)min(e, last) * rbs + c);
      }
    }
  }
  {   // object root rows
#pragma unroll
    for (int i = 0; i < (kTile * 13 + NT - 1) / NT; ++i) {
      const int j = tid + NT * i;
      if (j < kTile * 13) {
        const int e = j / 13, c = j - 13 * e, env = env0 + min(e, last);
        const size_t row = p.root_regular ? (size_t)env * p.actors_per_env + p.object_slot : (size_t)__ldg(b.object_actor + env);
        cp_async4(s.obj + j, b.root_state + row * 13 + c);
      }
    }
  }
  {   // fingertip rows: pos + quat (7 of 13 floats)
#pragma unroll
    for (int i = 0; i < (kTile * kT1 * 7 + NT - 1) / NT; ++i) {
      const int j = tid + NT * i;
      if (j < kTile * kT1 * 7) {
        const int e = j / (kT1 * 7), r = j - (kT1 * 7) * e, t = r / 7, c = r - 7 * t;
        cp_async4(s.tip + e * S::kTipW + r, rb0 + (size_t)min(e, last) * rbs + (size_t)p.tip_body[t] * 13 + c);
      }
    }
  }
  }   // !SOA
  if (nvalid == kTile) {   // goal rows: the tile's 32 x 13 floats are one contiguous, 16-byte aligned run (whole lines are fetched anyway)
    if (tid < kTile * 13 / 4) cp_async16(s.goal + 4 * tid, b.goal_states + (size_t)env0 * 13 + 4 * tid);
  } else {                 // ragged last tile: pose columns only
    for (int j = tid; j < kTile * 7; j += NT) {
      const int e = j / 7, c = j - 7 * e;
      cp_async4(s.goal + e * 13 + c, b.goal_states + (size_t)(env0 + min(e, last)) * 13 + c);
    }
  }
  const float* kp_src = b.object_keypoint_offsets + (size_t)env0 * 3 * K;
  const float* sc_src = b.object_scales + (size_t)env0 * 3;
  const float* dof_src = b.dof_state + (size_t)env0 * (2 * kD1);
  if (bulk && HA_BULK_IMAGES) {   // the three contiguous images: one bulk copy each
    if (tid == 32) bulk_g2s(s.kp, kp_src, kTile * 3 * K * 4, bar);
    if (tid == 33) bulk_g2s(s.scales, sc_src, kTile * 3 * 4, bar);
    if (tid == 34) bulk_g2s(s.dof, dof_src, kTile * 2 * kD1 * 4, bar);
  } else if (nvalid == kTile) {   // contiguous tiles: 16-byte copies (tile bases are 16-byte aligned: env0 % 32 == 0)
    constexpr int kKpVec = kTile * 3 * K / 4, kScVec = kTile * 3 / 4, kDofVec = kTile * 2 * kD1 / 4;
    if (tid < kKpVec) cp_async16(reinterpret_cast<float4*>(s.kp) + tid, reinterpret_cast<const float4*>(kp_src) + tid);
    if (tid >= NT - kScVec) {
      const int q = tid - (NT - kScVec);
      cp_async16(reinterpret_cast<float4*>(s.scales) + q, reinterpret_cast<const float4*>(sc_src) + q);
    }
#pragma unroll
    for (int i = 0; i < (kDofVec + NT - 1) / NT; ++i) {
      const int q = tid + NT * i;
      if (q < kDofVec) cp_async16(reinterpret_cast<float4*>(s.dof) + q, reinterpret_cast<const float4*>(dof_src) + q);
    }
  } else {                 // ragged last tile: element copies, nothing is read past the end of the arrays
    for (int j = tid; j < nvalid * 3 * K; j += NT) cp_async4(s.kp + j, kp_src + j);
    for (int j = tid; j < nvalid * 3; j += NT) cp_async4(s.scales + j, sc_src + j);
    for (int j = tid; j < nvalid * 2 * kD1; j += NT) cp_async4(s.dof + j, dof_src + j);
  }
}

// per-env state of one tile, loaded by warp 3 (lane = env: every request is one coalesced 128 / 256-byte run)
struct KukaPostEnvState {
  long long progress, reset, randomize;
  float succ, ckp, furthest, initz, prev_ep, old_rew_obs;
  int near;
  unsigned char lifted;
  float4 ctip, coef;
};

template <int OBS, int REW_SLOT>
__device__ __forceinline__ void kuka_post_load_env(const HaKukaParams& p, const HaKukaBuffers& b, const int env, const int mode,
                                                   KukaPostEnvState& r) {
  r.progress = b.progress_buf[env];
  r.reset = b.reset_buf[env];
  r.succ = b.successes[env];
  r.ckp = b.closest_keypoint_max_dist[env];
  r.furthest = b.furthest_hand_dist[env];
  r.near = b.near_goal_steps[env];
  r.lifted = b.lifted_object[env];
  r.initz = b.object_init_z ? ldg(b.object_init_z + env) : ldg(b.object_init_state + (size_t)env * 13 + 2);
  r.prev_ep = b.prev_episode_successes[env];
  r.ctip = *reinterpret_cast<const float4*>(b.closest_fingertip_dist + (size_t)env * 4);
  r.coef = __ldg(reinterpret_cast<const float4*>(b.finger_rew_coeffs + (size_t)env * 4));
  r.old_rew_obs = 0.f;
  r.randomize = 0;
  if (mode == kPostObserveOnly) {
    r.old_rew_obs = b.obs_buf[(size_t)env * OBS + REW_SLOT];
  } else {
    if (b.randomize_buf) r.randomize = b.randomize_buf[env];
  }
}

template <int K, bool SOA, int NS>
__global__ void __launch_bounds__(kPostThreads, NS == 1 ? HA_POST_NS1_CTAS : 4)
kuka_post_kernel(const __grid_constant__ HaKukaParams p, const __grid_constant__ HaKukaBuffers b, const int mode) {
  using SM = KukaPostSmem<K, NS>;
  using ST = KukaPostStage<K>;
  constexpr int OBS = SM::kObs, TIPW = ST::kTipW;
  constexpr KukaLayout L = kuka_layout(kD1, kT1, K);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  SM& sm = *reinterpret_cast<SM*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31;
  // roles rotate with the CTA index: warp w of every CTA sits on scheduler w % 4, so a fixed role <-> warp
  // mapping would pile all reward warps of an SM onto one scheduler
  const int warp = ((tid >> 5) + blockIdx.x) & 3;
  const int n = p.n_envs;
  const int n_tiles = (n + kTile - 1) / kTile;
  const int stride = gridDim.x;
  int tile = blockIdx.x;

  HA_TRACE_BEGIN(1);
  pdl_launch_dependents();      // the next kernel in the stream may start launching as our CTAs retire
  pdl_wait();                   // everything below reads what the previous kernel wrote
  float keypoint_tol = 0.f, tobj_scale = 0.f, tobj_offset = 0.f;
  if (warp == 3) {
    keypoint_tol = b.dynamic->keypoint_success_tolerance;
    tobj_scale = b.dynamic->true_objective_scale;
    tobj_offset = b.dynamic->true_objective_offset;
  }
  if (blockIdx.x == 0 && tid == 0 && mode == kPostFull) {
    // step bookkeeping that needs no ordering against the rest of the launch: the reset / index counters of the pre-step are
    // consumed by now (cleared for the next pre-step), the fast-mode RNG ticks once per step
    b.reset_counts[0] = 0; b.reset_counts[1] = 0; b.reset_counts[2] = 0; b.reset_counts[3] = 0;
    b.dynamic->rng_step = b.dynamic->rng_step + 1;
  }
  KukaPostAcc acc;
  if constexpr (SOA) {
    if (HA_BULK_ROWS || HA_BULK_IMAGES) {
      if (tid == 0) { mbar_init(&sm.mbar[0], 1); mbar_init(&sm.mbar[1], 1); mbar_fence_init(); }
    }
    // plane-row sources of the staged rows, decoded once per CTA (the per-tile issue then costs one shared load per copy)
    if (tid < 13 + 13 + kT1 * 7) {
      const size_t nn = (size_t)n;
      const int r = tid;
      const float* src;
      if (r < 13) src = b.rigid_body_state + ((size_t)p.palm_body[0] * 13 + r) * nn;
      else if (r < 26) src = b.root_state + ((size_t)p.object_slot * 13 + (r - 13)) * nn;
      else { const int rr = r - 26, t = rr / 7, c = rr - 7 * t; src = b.rigid_body_state + ((size_t)p.tip_body[t] * 13 + c) * nn; }
      sm.rowp[r] = src;
    }
    __syncthreads();
  }

  if (tile < n_tiles) kuka_post_issue<K, SOA>(p, b, sm.st[0], tile * kTile, min(kTile, n - tile * kTile), tid, &sm.mbar[0], sm.rowp);
  cp_async_commit();
  HA_TRACE_POINT(1, 1);         // prologue done, first tile's copies issued

  for (int it = 0; tile < n_tiles; ++it, tile += stride) {
    ST& S = sm.st[NS == 2 ? (it & 1) : 0];
    const int env0 = tile * kTile;
    const int nvalid = min(kTile, n - env0);
    const bool valid = lane < nvalid;
    const int env = env0 + min(lane, nvalid - 1);       // idle lanes of a ragged tile re-read a valid env; stores are predicated
    const int next = tile + stride;
    if constexpr (NS == 2 && !HA_LATE_PREFETCH) {   // next tile's copies go in flight before this tile is touched
      if (next < n_tiles) kuka_post_issue<K, SOA>(p, b, sm.st[(it + 1) & 1], next * kTile, min(kTile, n - next * kTile), tid, &sm.mbar[(it + 1) & 1], sm.rowp);
      cp_async_commit();
    }
    KukaPostEnvState r;
    if (warp == 3) kuka_post_load_env<OBS, L.reward>(p, b, env, mode, r);
    // per-episode reward sums (rewards_episode[k] += term, :871-874, :923-926): kept by roles 0 and 1, half of the terms each -- their loads
    // fly with the tile, the sums leave after B3 with the terms the state warp parked in shared memory (its B2 -> B3 chain is the
    // CTA's critical path; this takes 32 memory instructions and twelve registers out of it)
    constexpr int kHalfTerms = HA_KUKA_N_EPISODE_TERMS / 2;
    static_assert(HA_KUKA_N_EPISODE_TERMS == 2 * kHalfTerms, "two roles x half of the terms");
    float accv[kHalfTerms];
#pragma unroll
    for (int j = 0; j < kHalfTerms; ++j) accv[j] = 0.f;
    if (warp < 2 && mode != kPostObserveOnly) {
#pragma unroll
      for (int j = 0; j < kHalfTerms; ++j) accv[j] = b.rewards_episode[(size_t)(kHalfTerms * warp + j) * n + env];
    }
    if (HA_BULK_STORE && tid == 0) bulk_wait_read0();   // the previous tile's bulk store has read the obs tile
    cp_async_wait<(NS == 2 && !HA_LATE_PREFETCH) ? 1 : 0>();
    if (kuka_post_bulk_tile(SOA, n, env0)) mbar_wait(&sm.mbar[NS == 2 ? (it & 1) : 0], NS == 2 ? (it >> 1) & 1 : it & 1);
    __syncthreads();                                    // B1: this tile's inputs are in shared memory
    HA_TRACE_POINT(1, it == 0 ? 2 : (it == 1 ? 10 : 13));
    if constexpr (NS == 2 && HA_LATE_PREFETCH) {
      // The next tile's reads start now, so they share the memory system with this tile's stores instead of with this tile's own
      // reads: issued up front, both stages of every CTA land together (the whole input, ~3.5 us) and only then do the stores start.
      if (next < n_tiles) kuka_post_issue<K, SOA>(p, b, sm.st[(it + 1) & 1], next * kTile, min(kTile, n - next * kTile), tid, &sm.mbar[(it + 1) & 1], sm.rowp);
      cp_async_commit();
    }

    float* o = sm.obs + lane * OBS;
    const float clamp_c = p.clamp_abs_obs;                // clamp_obs (:1174-1176) / vec_task.py:432, applied where a value is produced
    const bool do_clamp = clamp_c > 0.0f;
    auto W = [&](int i, float v) { o[i] = (HA_BULK_STORE && do_clamp) ? fminf(fmaxf(v, -clamp_c), clamp_c) : v; };   // else: clamped by the store loop
    float obj_z = 0.f;
    // staged rows: [e][13] / [e][29] images of the AoS rows, or component-major [c][32] planes (SOA)
    auto PALM = [&](int c) { return SOA ? S.palm[c * kTile + lane] : S.palm[lane * 13 + c]; };
    auto OBJ = [&](int c) { return SOA ? S.obj[c * kTile + lane] : S.obj[lane * 13 + c]; };
    auto TIP = [&](int t, int c) { return SOA ? S.tip[(t * 7 + c) * kTile + lane] : S.tip[lane * TIPW + t * 7 + c]; };
    if (warp == 0) {
      // palm_center_pos = palm_pos + quat_rotate(palm_rot, palm_offset)   (:1013-1017); fingertips (:1033-1052)
      const f3 pc = f3{PALM(0), PALM(1), PALM(2)} + quat_rotate(f4{PALM(3), PALM(4), PALM(5), PALM(6)}, f3{p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]});
      const f3 obj_pos = {OBJ(0), OBJ(1), OBJ(2)};
      W(L.palm_pos + 0, pc.x); W(L.palm_pos + 1, pc.y); W(L.palm_pos + 2, pc.z);
#pragma unroll
      for (int t = 0; t < kT1; ++t) {
        const f3 to = f3{TIP(t, 0), TIP(t, 1), TIP(t, 2)} +
                      quat_rotate(f4{TIP(t, 3), TIP(t, 4), TIP(t, 5), TIP(t, 6)}, f3{p.tip_offset[t][0], p.tip_offset[t][1], p.tip_offset[t][2]});
        const f3 rel_obj = to - obj_pos, rel_palm = to - pc;
        sm.xch[t * kTile + lane] = norm3(rel_obj);
        W(L.tip_rel_palm + 3 * t + 0, rel_palm.x);
        W(L.tip_rel_palm + 3 * t + 1, rel_palm.y);
        W(L.tip_rel_palm + 3 * t + 2, rel_palm.z);
      }
    } else if (warp == 1) {
      // keypoints (:1057-1073)
      const float* gl = S.goal + lane * 13;
      const f3 pc = f3{PALM(0), PALM(1), PALM(2)} + quat_rotate(f4{PALM(3), PALM(4), PALM(5), PALM(6)}, f3{p.palm_offset[0], p.palm_offset[1], p.palm_offset[2]});
      const f3 obj_pos = {OBJ(0), OBJ(1), OBJ(2)}, goal_pos = {gl[0], gl[1], gl[2]};
      const f4 obj_rot = {OBJ(3), OBJ(4), OBJ(5), OBJ(6)}, goal_rot = {gl[3], gl[4], gl[5], gl[6]};
      float kpv[3 * K];
      if constexpr (K == 4) {
        const float4* k4 = reinterpret_cast<const float4*>(S.kp) + lane * 3;
#pragma unroll
        for (int i = 0; i < 3; ++i) { const float4 v = k4[i]; kpv[4 * i] = v.x; kpv[4 * i + 1] = v.y; kpv[4 * i + 2] = v.z; kpv[4 * i + 3] = v.w; }
      } else {
#pragma unroll
        for (int i = 0; i < 3 * K; ++i) kpv[i] = S.kp[lane * 3 * K + i];
      }
      float kp_max = 0.0f;
#pragma unroll
      for (int k = 0; k < K; ++k) {
        const f3 off = {kpv[3 * k], kpv[3 * k + 1], kpv[3 * k + 2]};
        const f3 okp = obj_pos + quat_rotate(obj_rot, off);
        const f3 gkp = goal_pos + quat_rotate(goal_rot, off);
        const f3 rel_goal = okp - gkp, rel_palm = okp - pc;
        const float dist = norm3(rel_goal);
        kp_max = (k == 0) ? dist : fmaxf(kp_max, dist);
        W(L.kp_rel_palm + 3 * k + 0, rel_palm.x);
        W(L.kp_rel_palm + 3 * k + 1, rel_palm.y);
        W(L.kp_rel_palm + 3 * k + 2, rel_palm.z);
        W(L.kp_rel_goal + 3 * k + 0, rel_goal.x);
        W(L.kp_rel_goal + 3 * k + 1, rel_goal.y);
        W(L.kp_rel_goal + 3 * k + 2, rel_goal.z);
      }
      sm.xch[4 * kTile + lane] = kp_max;
    } else if (warp == 2) {
      // unscale(dof_pos) (:1095-1100), raw dof_vel (:1104), _action_penalties sums (:830-839)
      const float2* dl = reinterpret_cast<const float2*>(S.dof) + lane * kD1;
      float sa = 0.f, sh = 0.f;
#pragma unroll
      for (int d = 0; d < kD1; ++d) {
        const float2 pv = dl[d];
        const float lo = p.dof_lower[d], hi = p.dof_upper[d];
        W(L.dof_pos + d, div_by_const(2.0f * pv.x - hi - lo, hi - lo, p.dof_range_rcp[d]));
        W(L.dof_vel + d, pv.y);
        if (d < 7) sa += fabsf(pv.y); else sh += fabsf(pv.y);
      }
      sm.xch[5 * kTile + lane] = sa;
      sm.xch[6 * kTile + lane] = sh;
    } else {
      // palm rot + vel (:1112-1117), object rot + vel (:1120), object scales (:1143)
      obj_z = OBJ(2);
#pragma unroll
      for (int c = 0; c < 10; ++c) { W(L.palm_rot_vel + c, PALM(3 + c)); W(L.obj_rot_vel + c, OBJ(3 + c)); }
#pragma unroll
      for (int c = 0; c < 3; ++c) W(L.scales + c, S.scales[lane * 3 + c]);
    }
    HA_TRACE_POINT_IF(1, 3 + warp, it == 0 && lane == 0);   // role `warp` of tile 0 done
    __syncthreads();                                    // B2: distances, kp_max and |vel| sums are visible to warp 3
    HA_TRACE_POINT(1, it == 0 ? 7 : (it == 1 ? 11 : 13));

    // Observation tile store, split in two: every 128-bit word that holds none of the columns the state warp still has to write
    // (closest_kp .. reward, the last 9 of a row) is complete now and leaves through the three other warps while the state warp
    // runs the reward chain; the 3 remaining words per row follow after B3.
    constexpr int kLate0 = HA_EARLY_STORE ? L.closest_kp - 3 : 0;              // first column offset of a word that overlaps the late columns
    const bool clamp_here = do_clamp && !HA_BULK_STORE;
    float4* obs_dst = reinterpret_cast<float4*>(b.obs_buf + (size_t)env0 * OBS);   // 16-byte aligned: env0 % 32 == 0
    const float4* obs_src = reinterpret_cast<const float4*>(sm.obs);
    const int obs_total = nvalid * OBS, obs_nvec = obs_total >> 2;
    auto store_word = [&](int q) {
      float4 v = obs_src[q];
      if (clamp_here) {
        v.x = fminf(fmaxf(v.x, -clamp_c), clamp_c); v.y = fminf(fmaxf(v.y, -clamp_c), clamp_c);
        v.z = fminf(fmaxf(v.z, -clamp_c), clamp_c); v.w = fminf(fmaxf(v.w, -clamp_c), clamp_c);
      }
      obs_dst[q] = v;
    };
    if (HA_EARLY_STORE && !HA_BULK_STORE && warp != 3) {
      constexpr int kStep = kPostThreads - 32, kColStep = (4 * kStep) % OBS;   // column offset of a thread's next word
      int q = warp * 32 + lane, c = (4 * q) % OBS;
#pragma unroll 2
      for (; q < obs_nvec; q += kStep) {
        if (c < kLate0) store_word(q);
        c += kColStep;
        c -= (c >= OBS) ? OBS : 0;
      }
    }

    if (warp == 3) {
      float cur[kT1];
#pragma unroll
      for (int t = 0; t < kT1; ++t) cur[t] = sm.xch[t * kTile + lane];
      const float kp_max = sm.xch[4 * kTile + lane], sa = sm.xch[5 * kTile + lane], sh = sm.xch[6 * kTile + lane];
      float ctip[4] = {r.ctip.x, r.ctip.y, r.ctip.z, r.ctip.w};
      const float coef[4] = {r.coef.x, r.coef.y, r.coef.z, r.coef.w};
      // lazy initialisation of the per-episode closest / furthest distances (:1044-1049, :1077-1079)
#pragma unroll
      for (int i = 0; i < 4; ++i) ctip[i] = (ctip[i] < 0.0f) ? cur[i] : ctip[i];
      float furthest = (r.furthest < 0.0f) ? cur[0] : r.furthest;
      float ckp = (r.ckp < 0.0f) ? kp_max : r.ckp;
      const long long progress = r.progress + (mode == kPostFull ? 1 : 0);   // progress_buf += 1 (:1429)
      // observation slots that see pre-update state (ordering: SURVEY 8a)
      W(L.closest_kp, ckp);
#pragma unroll
      for (int i = 0; i < 4; ++i) W(L.closest_tip + i, ctip[i]);
      W(L.lifted, r.lifted ? 1.0f : 0.0f);
      W(L.progress, logf((float)progress / 10.0f + 1.0f));
      W(L.successes, logf(r.succ + 1.0f));
      if (mode == kPostObserveOnly) {
        W(L.reward, r.old_rew_obs);
        if (valid) {
          b.closest_keypoint_max_dist[env] = ckp;
          b.furthest_hand_dist[env] = furthest;
          *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
        }
      } else {
        // _lifting_reward (:788-812)
        const float z_lift = 0.05f + obj_z - r.initz;
        float lifting_rew = clipf(z_lift, 0.0f, 0.5f);
        const bool lifted_prev = r.lifted != 0;
        const bool lifted = (z_lift > p.lifting_bonus_threshold) || lifted_prev;
        const float lift_bonus_rew = p.lifting_bonus * ((lifted && !lifted_prev) ? 1.0f : 0.0f);
        const float not_lifted = lifted ? 0.0f : 1.0f;
        lifting_rew *= not_lifted;
        // _distance_delta_rewards (:759-786)
        float ft = 0.f;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float delta = ctip[i] - cur[i];
          ctip[i] = fminf(ctip[i], cur[i]);
          ft += clipf(delta, 0.0f, 10.0f) * coef[i];
        }
        const float hand_delta = furthest - cur[0];
        furthest = fmaxf(furthest, cur[0]);
        ft *= not_lifted;
        float hp = clipf(hand_delta, -10.0f, 0.0f);
        hp *= not_lifted;
        hp *= (float)kT1;
        // _keypoint_reward (:814-828)
        float kp_rew = clipf(ckp - kp_max, 0.0f, 100.0f);
        ckp = fminf(ckp, kp_max);
        kp_rew *= lifted ? 1.0f : 0.0f;
        // compute_kuka_reward (:854-930)
        const bool near_goal = kp_max <= keypoint_tol;
        const int near_steps = r.near + (near_goal ? 1 : 0);
        const bool is_success = near_steps >= p.success_steps;
        const float succ = r.succ + (is_success ? 1.0f : 0.0f);
        float t[HA_KUKA_N_EPISODE_TERMS];
        t[HA_KUKA_RAW_FINGERTIP_DELTA] = ft; t[HA_KUKA_RAW_HAND_DELTA] = hp; t[HA_KUKA_RAW_LIFTING] = lifting_rew; t[HA_KUKA_RAW_KEYPOINT] = kp_rew;
        ft *= p.distance_delta_rew_scale;
        hp *= p.hand_delta_rew_scale;
        lifting_rew *= p.lifting_rew_scale;
        kp_rew *= p.keypoint_rew_scale;
        const float kuka_pen = -1.0f * (sa * p.kuka_actions_penalty_scale);
        const float allegro_pen = -1.0f * (sh * p.allegro_actions_penalty_scale);
        const float bonus_rew = (near_goal ? 1.0f : 0.0f) * p.bonus_per_step;
        const float reward = ft + hp + lifting_rew + lift_bonus_rew + kp_rew + kuka_pen + allegro_pen + bonus_rew;
        t[HA_KUKA_FINGERTIP_DELTA] = ft; t[HA_KUKA_HAND_DELTA] = hp; t[HA_KUKA_LIFTING] = lifting_rew; t[HA_KUKA_LIFT_BONUS] = lift_bonus_rew;
        t[HA_KUKA_KEYPOINT] = kp_rew; t[HA_KUKA_BONUS] = bonus_rew; t[HA_KUKA_KUKA_PENALTY] = kuka_pen; t[HA_KUKA_ALLEGRO_PENALTY] = allegro_pen;
        // _compute_resets (:841-849) + reorientation _extra_reset_rules (allegro_kuka_reorientation.py:97-102)
        long long resets = (obj_z < p.fall_z) ? 1 : r.reset;
        long long prog = progress;
        if (p.max_consecutive_successes > 0) {
          prog = is_success ? 0 : prog;
          resets = (succ >= (float)p.max_consecutive_successes) ? 1 : resets;
        }
        const bool timed_out = prog >= (long long)(p.max_episode_length - 1);
        resets = timed_out ? 1 : resets;
        if (p.subtask == HA_KUKA_REORIENTATION) {
          const float far = fmaxf(fmaxf(cur[0], cur[1]), fmaxf(cur[2], cur[3]));
          resets = (far > p.far_tip_dist) ? 1 : resets;
        }
        const float true_obj = succ * tobj_scale + tobj_offset;
        const bool timeout_flag = timed_out && (resets != 0);          // vec_task.py:424
        W(L.reward, reward * 0.01f);                                  // (:1437-1439)
        if (valid) {
          // per-env outputs (lane = env: coalesced)
          b.rew_buf[env] = reward;
          b.reset_buf[env] = resets;
          b.progress_buf[env] = prog;
          b.reset_goal_buf[env] = is_success ? 1 : 0;
          if (b.randomize_buf) b.randomize_buf[env] = r.randomize + 1;   // (:1430)
          b.timeout_buf[env] = timeout_flag ? 1 : 0;
          b.successes[env] = succ;
          b.true_objective[env] = true_obj;
          b.near_goal_steps[env] = near_steps;
          b.lifted_object[env] = lifted ? 1 : 0;
          b.closest_keypoint_max_dist[env] = ckp;
          b.furthest_hand_dist[env] = furthest;
          *reinterpret_cast<float4*>(b.closest_fingertip_dist + (size_t)env * 4) = make_float4(ctip[0], ctip[1], ctip[2], ctip[3]);
          // the terms of this step: parked for the other roles (rewards_episode, episode_cumulative; after B3)
#pragma unroll
          for (int k = 0; k < HA_KUKA_N_EPISODE_TERMS; ++k) sm.terms[k * kTile + lane] = t[k];
          // statistics partials (fixed point: order-independent sums)
          acc.rew += __float2ll_rn(reward * stat_scale(HA_KUKA_STAT_SUM_REWARD));
          acc.prev_ep += __float2ll_rn(r.prev_ep * stat_scale(HA_KUKA_STAT_SUM_PREV_EP_SUCCESSES));
          acc.tobj += __float2ll_rn(true_obj * stat_scale(HA_KUKA_STAT_SUM_TRUE_OBJECTIVE));
          acc.mx = fmaxf(acc.mx, true_obj);
          acc.neg_mn = fmaxf(acc.neg_mn, -true_obj);
          acc.n += 1; acc.resets += (resets != 0); acc.success += is_success; acc.timeout += timeout_flag; acc.lifted += lifted;
        }
      }
    }
    const bool bulk_store = HA_BULK_STORE && (nvalid == kTile);   // 14976 contiguous bytes, 16-byte aligned on both sides
    if (bulk_store) fence_proxy_async_smem();           // the tile's generic-proxy writes, before the async proxy reads them
    HA_TRACE_POINT_IF(1, 8, it == 0 && warp == 3 && lane == 0);   // reward chain of tile 0 done
    __syncthreads();                                    // B3: the observation tile is complete
    HA_TRACE_POINT(1, it == 0 ? 9 : (it == 1 ? 12 : 13));
    if (mode != kPostObserveOnly && valid) {
      if (warp < 2) {                                   // rewards_episode[k] += term (:871-874, :923-926)
#pragma unroll
        for (int j = 0; j < kHalfTerms; ++j) b.rewards_episode[(size_t)(kHalfTerms * warp + j) * n + env] = accv[j] + sm.terms[(kHalfTerms * warp + j) * kTile + lane];
      } else if (warp == 2) {                           // episode_cumulative[k] = term, in the reference's list order (:912-921)
        constexpr int cum_src[HA_KUKA_N_STEP_TERMS] = {HA_KUKA_FINGERTIP_DELTA, HA_KUKA_HAND_DELTA, HA_KUKA_LIFTING, HA_KUKA_LIFT_BONUS,
                                                       HA_KUKA_KEYPOINT, HA_KUKA_KUKA_PENALTY, HA_KUKA_ALLEGRO_PENALTY, HA_KUKA_BONUS};
#pragma unroll
        for (int k = 0; k < HA_KUKA_N_STEP_TERMS; ++k) b.episode_cumulative[(size_t)k * n + env] = sm.terms[cum_src[k] * kTile + lane];
      }
    }
    if (bulk_store) {
      if (tid == 0) {                                   // the whole [32, OBS] tile leaves as one bulk copy
        bulk_s2g(b.obs_buf + (size_t)env0 * OBS, sm.obs, kTile * OBS * 4);
        bulk_commit();
      }
    } else {
      if (!HA_EARLY_STORE) {
#pragma unroll 2
        for (int q = tid; q < obs_nvec; q += kPostThreads) store_word(q);
      } else if (tid < 3 * kTile) {   // the words that overlap the late columns: 3 per row
        const int r = tid / 3, j = tid - 3 * r;
        const int q = (OBS * r + kLate0 + 3) / 4 + j;
        if (q < obs_nvec) store_word(q);
      }
      const int f = (obs_nvec << 2) + tid;              // ragged last tile: up to 3 trailing floats
      if (f < obs_total) {
        float v = sm.obs[f];
        if (clamp_here) v = fminf(fmaxf(v, -clamp_c), clamp_c);
        b.obs_buf[(size_t)env0 * OBS + f] = v;
      }
    }
    if constexpr (NS == 1) {
      if (next < n_tiles) {
        __syncthreads();                                // every warp is done with the stage and the obs tile
        kuka_post_issue<K, SOA>(p, b, sm.st[0], next * kTile, min(kTile, n - next * kTile), tid, &sm.mbar[0], sm.rowp);
      }
      cp_async_commit();
    }
  }
  cp_async_wait<0>();
  // cross-rank statistics without a launch (HaKukaBuffers.xchg): the last CTA of the grid owns the fewest tiles and is done first;
  // after its tiles one of its warps publishes the statistics vector the preceding pre step folded (nobody writes stats[] during
  // this launch) and reduces the newest complete exchange.  Remote stores and local loads only, nothing waits for a peer.
  if (b.xchg != nullptr && mode == kPostFull && blockIdx.x == gridDim.x - 1 && (tid >> 5) == 0) {
    const HaStatsExchange& x = *reinterpret_cast<const HaStatsExchange*>(b.xchg);
    const long long calls = x.seq[2] + 1;
    __syncwarp();
    if (lane == 0) x.seq[2] = calls;
    if (x.interval <= 1 || calls % x.interval == 0) stats_exchange_warp(x, b.stats, b.xchg_reduced, lane);
  }
  if (HA_BULK_STORE && tid == 0) bulk_wait_read0();     // shared memory must outlive the last bulk store's reads
  if (warp == 3 && mode == kPostFull) kuka_post_store_partials(b, acc, lane);   // every CTA of the grid fills its slot
  HA_TRACE_POINT(1, 13);
  HA_TRACE_END(1);
}

}  // namespace ha
