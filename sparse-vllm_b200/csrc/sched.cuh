// Dynamic chunk scheduling of the persistent page-scoring kernel.
//
// Why: with one equal static cut per SM the same SMs are always served ~20 % slower by the memory
// system (profiles/r1_c_ncu_summary.md) and the fast ones idle.  Here the work of a launch is
// rows x chunks (a chunk = up to `cu` consecutive units of one row, sized to ~64 KB by the launcher)
// claimed with atomicAdd on the row's counter.  A CTA starts on its home rows (group = blockIdx %
// groups; rows group, group + groups, ...).  Its FIRST batch on its first home row is static (the
// l-th home CTA of the row owns chunks [l*m0, (l+1)*m0): no atomic round trip before the first byte
// moves); after that it claims guided batches: a fraction of its fair share of what is left, so
// batches are large while a row is far from done (the claim's round trip hides behind queued work)
// and shrink near the end (little is hoarded when the row runs dry).  Once its home rows are
// exhausted (or if it has none: gridDim is not a multiple of the groups) it steals a few chunks at a
// time from the row with the most chunks left.
//
// One scheduler warp per CTA claims and posts descriptors into a shared-memory queue (full/empty
// mbarrier pair per slot); every producer and consumer warp reads every descriptor in order.
#pragma once

#include "common.cuh"

namespace svb200 {

constexpr int kMaxChunkUnits = 8;  // units per claimed chunk at most (a producer lane looks up one unit's chunk id)
constexpr int kSchedQueue = 32;  // descriptor slots (sched_run's `depth` <= kSchedQueue bounds what is actually posted ahead)
constexpr int kMinBatch = 2;     // chunks per claim at least (an atomicAdd costs 3-4 us under load; one more is in flight)

struct ChunkDesc {
  int row;         // -1: end of work
  int unit_begin;  // first unit inside the row
  int n_units;     // 1..kMaxChunkUnits, | aux << 8 (aux: what reserve(row) returned for this visit of the row)
  int unit_base;   // CTA-local running unit index of the chunk's first unit (ring stage = index % S)
  __device__ __forceinline__ int units() const { return n_units & 0xff; }
  __device__ __forceinline__ int aux() const { return n_units >> 8; }
};

struct SchedSmem {
  ChunkDesc desc[kSchedQueue];
  uint64_t full[kSchedQueue];
  uint64_t empty[kSchedQueue];
};

// Global scratch of one launch family: ints [0] = finished CTAs, [1 + r] = next chunk of row r,
// [1 + rows + r] = units of row r done, [1 + 2 * rows + r] = row r selected (one-launch layer only).
// Zero-initialised once; the last CTA to finish zeroes it again.
__host__ __device__ inline int sched_scratch_ints(int rows) { return 1 + 3 * rows; }

__device__ __forceinline__ void sched_init(SchedSmem& q, int readers) {
  for (int s = 0; s < kSchedQueue; ++s) {
    mbar_init(&q.full[s], 1);
    mbar_init(&q.empty[s], readers);
  }
}

// Reader side (whole warp): descriptor n, released as soon as it is copied to registers.
__device__ __forceinline__ ChunkDesc sched_read(SchedSmem& q, int n) {
  const int slot = n % kSchedQueue;
  mbar_wait(&q.full[slot], (uint32_t)(n / kSchedQueue) & 1u);
  const ChunkDesc d = q.desc[slot];
  __syncwarp();
  if ((threadIdx.x & 31) == 0) mbar_arrive(&q.empty[slot]);
  return d;
}

__device__ __forceinline__ int ld_volatile_shared(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)) : "memory");
  return v;
}
__device__ __forceinline__ void st_volatile_shared(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

// Scheduler warp.  units_of_row(r) -> units of row r (uniform across the launch).
// reserve(row) is called once before a row is drained: < 0 = this CTA may not work on the row,
// else the value travels in the row's descriptors (aux).
struct ReserveNone {
  __device__ __forceinline__ int operator()(int) const { return 0; }
};
// share_div == 0: the first home row of every CTA is cut statically and evenly IN UNITS among its `share` home
// CTAs (descriptors of up to cu units, no counter traffic at all); nothing of those rows is left to claim or steal.
// For launches that are over before claims could rebalance anything (see the launcher).
template <class UnitsOfRow, class Reserve = ReserveNone>
__device__ void sched_run(SchedSmem& q, int* next_chunk /* global [rows] */, int rows, int groups, int cu,
                          int min_batch, int share_div, int depth, UnitsOfRow units_of_row,
                          Reserve reserve = Reserve(), unsigned long long* t_last_claim = nullptr,
                          int tail_chunks = 0) {
  bool even = share_div == 0;
  if (even) {
    // an even cut gives every row the same number of CTAs: only for rows of (nearly) equal length; ragged batches
    // need the claims and the stealing (every CTA reaches the same verdict from the same lengths)
    int64_t total = 0;
    int longest = 0;
    for (int r = 0; r < rows; ++r) {
      const int nu = units_of_row(r);
      total += nu;
      longest = max(longest, nu);
    }
    even = (int64_t)longest * rows * 4 <= total * 5;
    share_div = even ? 1 : 2;
  }
  const int lane = threadIdx.x & 31;
  int n = 0, ubase = 0, aux = 0;
  auto post = [&](int row, int begin, int cnt) {
    const int slot = n % kSchedQueue;
    // at most `depth` (<= kSchedQueue) descriptors are posted ahead of the slowest reader: what is
    // posted is committed to this CTA and cannot be stolen
    if (n >= depth)
      mbar_wait(&q.empty[(n - depth) % kSchedQueue], (uint32_t)((n - depth) / kSchedQueue) & 1u);
    if (lane == 0) {
      q.desc[slot] = ChunkDesc{row, begin, cnt | (aux << 8), ubase};
      mbar_arrive(&q.full[slot]);
    }
    __syncwarp();
    ubase += cnt;
    ++n;
  };
  // claims chunks of `row` until it is exhausted; the next claim is in flight while the current
  // batch is posted
  const int share = max(1, (int)gridDim.x / groups);  // CTAs that start on the same rows
  constexpr int kStealVisit = 12;  // chunks a thief takes per visit of a row before it looks again
  auto chunks_of = [&](int row) { return (units_of_row(row) + cu - 1) / cu; };
  auto guided_batch = [&](int left) { return max(min_batch, left / (share_div * share)); };
  // chunks of `row` that are handed out statically (first batch of the CTAs whose first home row it is);
  // the row's counter counts the chunks claimed BEYOND them.  Every CTA computes the same value.
  // (a static piece costs no claim round trip, so it may be a single chunk: small launches then spread over
  // every CTA instead of min_batch chunks going to each of a few)
  auto static_piece = [&](int d) { return max(1, d / (share_div * share)); };
  auto static_chunks = [&](int row) {
    if (row >= groups || share * groups > (int)gridDim.x) return 0;
    const int d = chunks_of(row);
    return even ? d : min(d, share * static_piece(d));
  };
  auto drain = [&](int row, bool guided, bool first_home) -> bool {  // false: reserve() refused the row
    const int nu = units_of_row(row);
    const int d = (nu + cu - 1) / cu;
    if (d == 0) return true;
    const int base = static_chunks(row);
    if (!guided && base + __ldcg(next_chunk + row) >= d) return true;  // thief, nothing left: do not reserve
    aux = reserve(row);
    if (aux < 0) return false;
    if (even && base == d) {
      // evenly cut row: only its home CTAs get here (for a thief nothing is left), each posts its own units
      if (first_home) {
        const int l = (int)(blockIdx.x / (unsigned)groups);
        const int u_end = (int)((int64_t)nu * (l + 1) / share);
        if (t_last_claim && lane == 0) *t_last_claim = global_timer_ns();
        for (int ub = (int)((int64_t)nu * l / share); ub < u_end; ub += cu) post(row, ub, min(cu, u_end - ub));
      }
      return true;
    }
    // the row's tail (its last tail_chunks chunks): one chunk per claim, one unit per descriptor -- what a CTA holds
    // cannot be stolen, so it should hold little when the row runs dry
    auto batch = [&](int left) { return left <= tail_chunks ? 1 : (guided ? guided_batch(left) : min_batch); };
    auto post_chunk = [&](int c) {
      const int ub = c * cu, ue = min(ub + cu, nu);
      if (d - c <= tail_chunks)
        for (int u = ub; u < ue; ++u) post(row, u, 1);
      else
        post(row, ub, ue - ub);
    };
    int m = first_home ? static_piece(d) : batch(d), j = 0, taken = 0;
    if (first_home) {
      j = (int)(blockIdx.x / (unsigned)groups) * m;  // static: no round trip before the first copy
      if (j >= base) first_home = false, m = batch(d);  // (the row is shorter than the static pieces)
    }
    if (!first_home) {
      if (lane == 0) j = base + atomicAdd(next_chunk + row, m);
      j = __shfl_sync(0xffffffffu, j, 0);
    }
    while (j < d) {
      if (t_last_claim && lane == 0) *t_last_claim = global_timer_ns();
      // a static piece ends where the static region does; what is left for claims starts behind that region
      const int end = min(j + m, first_home ? base : d);
      const int claimed_to = first_home ? base : end;
      first_home = false;
      taken += end - j;
      const int left = (guided || taken < kStealVisit) ? d - claimed_to : 0;
      int m2 = 0, j2 = 0;
      if (left > 0) {
        m2 = batch(left);
        if (lane == 0) j2 = base + atomicAdd(next_chunk + row, m2);
      }
      for (int c = j; c < end; ++c) post_chunk(c);
      if (left <= 0) return true;
      j = __shfl_sync(0xffffffffu, j2, 0);
      m = m2;
    }
    return true;
  };
  // share * groups CTAs own home rows; the others help wherever most is left from the start
  if ((int)blockIdx.x < share * groups) {
    const int grp = (int)(blockIdx.x % (unsigned)groups);
    for (int row = grp; row < rows; row += groups) drain(row, true, row == grp);
  }
  // steal: the row with the most chunks left (counters only grow: a stale read can only show too much)
  while (true) {
    int best = 0, best_row = -1;
    for (int base = 0; base < rows; base += 32) {
      const int r = base + lane;
      if (r < rows) {
        const int rr = (r + (int)blockIdx.x) % rows;
        const int nu = units_of_row(rr);
        const int left = (nu + cu - 1) / cu - static_chunks(rr) - __ldcg(next_chunk + rr);
        if (left > best) best = left, best_row = rr;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const int ob = __shfl_xor_sync(0xffffffffu, best, o);
      const int orow = __shfl_xor_sync(0xffffffffu, best_row, o);
      if (ob > best || (ob == best && orow > best_row)) best = ob, best_row = orow;
    }
    if (best <= 0) break;
    if (!drain(best_row, false, false)) break;  // contributor slots of that row are used up: stop stealing
  }
  post(-1, 0, 0);
}

// Last CTA out zeroes the scratch for the next launch / graph replay.  Called by one thread after
// all of the CTA's global side effects (and after its scheduler warp posted the end descriptor).
__device__ __forceinline__ void sched_finish(int* scratch, int rows, int extra = 0) {
  __threadfence();
  const int f = atomicAdd(scratch, 1);
  if (f == (int)gridDim.x - 1) {
    for (int i = 0; i < sched_scratch_ints(rows) + extra; ++i) scratch[i] = 0;
  }
}

}  // namespace svb200
