// Quest kernels for sm_100a: page min/max bounds, fused gather-free page scoring,
// radix top-k page selection + slot packing.
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/engine/cache_manager/quest.py):
//   bounds  : on_kv_stored 1608-1716, on_forward_end 1719-1771 (HBM layout: bounds_layout.cuh)
//   scoring : _score_pages_batched 1773-1802 (three bf16 roundings per head score)
//   select  : _build_decode_view_static 1833-1913
#include "bounds_layout.cuh"
#include <type_traits>

#include "common.cuh"
#include "decode_attn.cuh"
#include "launch.h"
#include "sched.cuh"
#include "store_kv.cuh"

namespace svb200 {

// =========================================================================
// Page bounds (layout: bounds_layout.cuh).  plane 0 = max, plane 1 = min.
// =========================================================================
__device__ __forceinline__ void page_minmax_vec(const __nv_bfloat16* page_base, int64_t row_elems,
                                                uint4& mx, uint4& mn) {
  // page_base points at this thread's 8-element column group of row 0.
  uint4 rows[kPage];
#pragma unroll
  for (int r = 0; r < kPage; ++r) rows[r] = ldg_stream_16(page_base + r * row_elems);
  mx = rows[0];
  mn = rows[0];
  __nv_bfloat162* mxp = reinterpret_cast<__nv_bfloat162*>(&mx);
  __nv_bfloat162* mnp = reinterpret_cast<__nv_bfloat162*>(&mn);
#pragma unroll
  for (int r = 1; r < kPage; ++r) {
    const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&rows[r]);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      mxp[i] = __hmax2(mxp[i], x[i]);
      mnp[i] = __hmin2(mnp[i], x[i]);
    }
  }
}

// One warp per token that completes a page (slot % 16 == 15).
__global__ void __launch_bounds__(256) quest_page_minmax_kernel(
    const __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ bounds,
    const int32_t* __restrict__ page_rec, const int32_t* __restrict__ slot_mapping, int n_tokens, int hkv,
    int64_t num_slots, int64_t num_recs) {
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const int vecs = hkv * (kHeadDim / 8);
  const int hs = bounds_head_split(hkv);
  for (int tok = blockIdx.x * warps_per_block + (threadIdx.x >> 5); tok < n_tokens;
       tok += gridDim.x * warps_per_block) {
    const int slot = slot_mapping[tok];
    if (slot < 0 || (slot & (kPage - 1)) != kPage - 1 || slot >= num_slots) continue;
    const int64_t page = slot >> 4;
    const int rec = page_rec[page];
    if (rec < 0 || rec >= num_recs) continue;
    const __nv_bfloat16* page_base = k_cache + page * kPage * row_elems;
    for (int v = lane; v < vecs; v += 32) {
      uint4 mx, mn;
      page_minmax_vec(page_base + v * 8, row_elems, mx, mn);
      *reinterpret_cast<uint4*>(bounds + bounds_seg_offset(rec, 0, v >> 4, v & 15, hkv, hs)) = mx;
      *reinterpret_cast<uint4*>(bounds + bounds_seg_offset(rec, 1, v >> 4, v & 15, hkv, hs)) = mn;
    }
  }
}

// grid (batch, layers): rebuild the bounds of the page a decode token just completed.
__global__ void __launch_bounds__(128) quest_decode_page_complete_kernel(
    const __nv_bfloat16* __restrict__ k_cache_all, __nv_bfloat16* __restrict__ bounds_all,
    const int32_t* __restrict__ page_rec, const int32_t* __restrict__ page_table, int64_t pt_stride,
    const int32_t* __restrict__ req_rows, const int32_t* __restrict__ seq_lens, int hkv,
    int64_t num_slots, int64_t num_pages, int64_t num_recs) {
  pdl_wait();
  const int b = blockIdx.x, layer = blockIdx.y;
  const int len = seq_lens[b];
  if (len <= 0 || (len & (kPage - 1)) != 0) return;
  const int64_t row = req_rows[b];
  const int slot_page = page_table[row * pt_stride + (len / kPage - 1)];
  if (slot_page < 0 || slot_page >= num_pages) return;
  const int rec = page_rec[slot_page];
  if (rec < 0 || rec >= num_recs) return;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const int hs = bounds_head_split(hkv);
  const __nv_bfloat16* page_base =
      k_cache_all + ((int64_t)layer * num_slots + (int64_t)slot_page * kPage) * row_elems;
  __nv_bfloat16* out = bounds_all + (int64_t)layer * num_recs * 2 * row_elems;
  const int vecs = hkv * (kHeadDim / 8);
  for (int v = threadIdx.x; v < vecs; v += blockDim.x) {
    uint4 mx, mn;
    page_minmax_vec(page_base + v * 8, row_elems, mx, mn);
    *reinterpret_cast<uint4*>(out + bounds_seg_offset(rec, 0, v >> 4, v & 15, hkv, hs)) = mx;
    *reinterpret_cast<uint4*>(out + bounds_seg_offset(rec, 1, v >> 4, v & 15, hkv, hs)) = mn;
  }
}

// =========================================================================
// Top-k selection + packing of one row (device routine shared by the standalone
// and the fused kernel).  Works for any block size that is a multiple of 32.
// =========================================================================
constexpr int kMaxPrevBudget = 2047;

struct SelectSmem {
  int wcnt[16][8];  // per search round: one count per warp of the group (<= 8 warps), summed by every thread
  int warp_scan[33];
  int sel[kMaxPrevBudget + 1];
  int sel_slot[kMaxPrevBudget + 1];
};

constexpr uint32_t kNegInfKey = 0x007fu;  // order key of bf16 -inf (0xff80)
// bf16 bit pattern -> 16-bit key that orders like the float value; never 0 (0 = "no such column").
__device__ __forceinline__ uint32_t bf16_order_key(uint32_t bits) {
  const uint32_t k = (bits & 0x8000u) ? (~bits & 0xffffu) : (bits | 0x8000u);
  return k ? k : 1u;
}

// Barrier over the first NTHR threads of the CTA (barrier 0 = __syncthreads when NTHR is the
// whole block; the fused scoring kernel keeps its producer warps out with barrier 1).
template <int BAR, int NTHR>
__device__ __forceinline__ void group_sync() {
  asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(NTHR) : "memory");
}

// Exclusive prefix of v over the group's threads; `total` receives the sum over the group.
template <int BAR, int NTHR, int TID0>
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_scan, int& total) {
  const int lane = threadIdx.x & 31, warp = (threadIdx.x - TID0) >> 5, nwarps = NTHR >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  group_sync<BAR, NTHR>();  // warp_scan may still be read from a previous call
  if (lane == 31) warp_scan[warp] = inc;
  group_sync<BAR, NTHR>();
  if (warp == 0) {
    const int w = lane < nwarps ? warp_scan[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += n;
    }
    warp_scan[lane] = winc - w;  // exclusive prefix of the warp totals
    if (lane == 31) warp_scan[32] = winc;
  }
  group_sync<BAR, NTHR>();
  total = warp_scan[32];
  return warp_scan[warp] + inc - v;
}

// scores_row: bf16 bits, 16-byte aligned, readable up to round_up(n_cols, 8).
// Columns >= n_valid count as -inf whatever the buffer holds (quest.py:1883-1885).
// The k-th largest 16-bit order key is found by a bitwise binary search (16 counting rounds over keys
// held in registers, one barrier each); a single ordered compaction then takes every key above it and
// the lowest-indexed ties.  Each thread owns a contiguous chunk of 8-score vectors.  CACHED keeps the
// chunk in registers so the row is read from L2 once (rows up to 4 * NTHR vectors = 8192 scores at
// 256 threads); longer rows re-read it every round.
constexpr int kSelectVecPerThread = 4;

template <int BAR, int NTHR, int CVPT, int TID0>  // CVPT: vectors a thread keeps in registers (0: none, long rows)
__device__ void quest_select_pack_impl(const uint16_t* __restrict__ scores_row, int n_cols, int n_valid,
                                       int k, const int32_t* __restrict__ pt_row, int ctx,
                                       int32_t* __restrict__ out_row, int32_t* __restrict__ len_out,
                                       int32_t* __restrict__ sel_pages_out, SelectSmem& sm,
                                       unsigned long long* tdbg = nullptr) {
  constexpr bool CACHED = CVPT > 0;
  constexpr int VPT = CACHED ? CVPT : 1;
  const unsigned long long tq0 = tdbg ? global_timer_ns() : 0;
  const int tid = threadIdx.x - TID0, lane = tid & 31;  // TID0: first thread of the group (a warp boundary)
  const int nvec = (n_cols + 7) >> 3;
  const uint4* vrow = reinterpret_cast<const uint4*>(scores_row);
  const int vec_per_thr = CACHED ? VPT : (nvec + NTHR - 1) / NTHR;
  const int v_begin = min(tid * vec_per_thr, nvec), v_end = min(v_begin + vec_per_thr, nvec);
  // one vector of 8 scores -> 4 words of packed order keys (low half = even column)
  auto to_keys = [&](int v, const uint4& x, uint32_t (&out)[4]) {
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int c0 = v * 8 + 2 * i, c1 = c0 + 1;
      const uint32_t lo = c0 >= n_cols ? 0u : (c0 < n_valid ? bf16_order_key(w[i] & 0xffffu) : kNegInfKey);
      const uint32_t hi = c1 >= n_cols ? 0u : (c1 < n_valid ? bf16_order_key(w[i] >> 16) : kNegInfKey);
      out[i] = lo | (hi << 16);
    }
  };
  uint32_t kw[VPT][4];
  if constexpr (CACHED) {
#pragma unroll
    for (int vv = 0; vv < VPT; ++vv) {
      if (v_begin + vv < v_end) {
        to_keys(v_begin + vv, __ldcg(vrow + v_begin + vv), kw[vv]);
      } else {
        kw[vv][0] = kw[vv][1] = kw[vv][2] = kw[vv][3] = 0u;
      }
    }
  }
  // body(first column of the word, packed key word), ascending column order
  auto for_each_word = [&](auto&& body) {
    if constexpr (CACHED) {
#pragma unroll
      for (int vv = 0; vv < VPT; ++vv)
#pragma unroll
        for (int i = 0; i < 4; ++i) body((v_begin + vv) * 8 + 2 * i, kw[vv][i]);
    } else {
      for (int v = v_begin; v < v_end; ++v) {
        uint32_t w[4];
        to_keys(v, __ldcg(vrow + v), w);
#pragma unroll
        for (int i = 0; i < 4; ++i) body(v * 8 + 2 * i, w[i]);
      }
    }
  };

  // ---- k-th largest key: largest T with count(key >= T) >= k
  const unsigned long long tq1 = tdbg ? global_timer_ns() : 0;
  static_assert(NTHR <= 256, "SelectSmem::wcnt holds one count per warp of at most 8 warps");
  constexpr int NWARP = NTHR / 32;
  group_sync<BAR, NTHR>();  // sm.wcnt may still be read by a previous call
  uint32_t thresh = 0;
  // total of a per-thread count over the group: one slot per (round, warp), no atomics, no reset
  auto group_total = [&](int bit, int c) {
    c = __reduce_add_sync(0xffffffffu, c);
    if (lane == 0) sm.wcnt[bit][tid >> 5] = c;
    group_sync<BAR, NTHR>();
    int total_c = 0;
#pragma unroll
    for (int w = 0; w < NWARP; ++w) total_c += sm.wcnt[bit][w];
    return total_c;
  };
  if constexpr (CACHED) {
    // Round 15 settles the top bit; the keys then shrink to 15 bits (keys the top bit already decided
    // become 0 = never counted, or 0x7fff = always counted), which lets one VIADDMNMX.S16x2.RELU
    // (max(min(key - cand + 1, 1), 0) per 16-bit half, no overflow with 15-bit operands) compare two
    // keys per instruction.  Measured on the 8192-column rows of cfg5 (8 warps, 32 keys per thread): compare +
    // add per key 4.0 us for the 16 rounds, this 2.9 us; two bits per round (three candidates, 9 barriers) 3.0 us
    // and a guarded subtraction on the integer pipe ((key | 0x8000) - cand keeps bit 15 iff key >= cand) 3.7 us:
    // the rounds are bound by the instructions of the 8 warps, not by their barriers.
    uint32_t acc = 0;
#pragma unroll
    for (int vv = 0; vv < VPT; ++vv)
#pragma unroll
      for (int i = 0; i < 4; ++i) acc += (kw[vv][i] >> 15) & 0x00010001u;
    const bool top = group_total(15, (int)((acc & 0xffffu) + (acc >> 16))) >= k;
    uint32_t sub[VPT][4];
#pragma unroll
    for (int vv = 0; vv < VPT; ++vv)
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t m = ((kw[vv][i] >> 15) & 0x00010001u) * 0x7fffu;
        sub[vv][i] = top ? (kw[vv][i] & m) : ((kw[vv][i] | m) & 0x7fff7fffu);
      }
    uint32_t t15 = 0;
    for (int bit = 14; bit >= 0; --bit) {
      const uint32_t cand = t15 | (1u << bit);
      const uint32_t addend = ((1u - cand) & 0xffffu) * 0x00010001u;
      uint32_t a0 = 0, a1 = 0;
#pragma unroll
      for (int vv = 0; vv < VPT; ++vv) {
        a0 += __viaddmin_s16x2_relu(sub[vv][0], addend, 0x00010001u) + __viaddmin_s16x2_relu(sub[vv][1], addend, 0x00010001u);
        a1 += __viaddmin_s16x2_relu(sub[vv][2], addend, 0x00010001u) + __viaddmin_s16x2_relu(sub[vv][3], addend, 0x00010001u);
      }
      a0 += a1;
      if (group_total(bit, (int)((a0 & 0xffffu) + (a0 >> 16))) >= k) t15 = cand;
    }
    thresh = (top ? 0x8000u : 0u) | t15;
  } else {
    for (int bit = 15; bit >= 0; --bit) {
      const uint32_t cand = thresh | (1u << bit), cand_hi = cand << 16;
      int c = 0;
      for_each_word([&](int, uint32_t w) {
        c += (w >= cand_hi) ? 1 : 0;
        c += ((w & 0xffffu) >= cand) ? 1 : 0;
      });
      if (group_total(bit, c) >= k) thresh = cand;
    }
  }

  // ---- ordered compaction: keys above thresh, then the lowest-indexed ties
  const unsigned long long tq2 = tdbg ? global_timer_ns() : 0;
  int cnt_gt = 0, cnt_eq = 0;
  for_each_word([&](int, uint32_t w) {
    const uint32_t lo = w & 0xffffu, hi = w >> 16;
    cnt_gt += (lo > thresh) + (hi > thresh);
    cnt_eq += (lo == thresh && lo) + (hi == thresh && hi);
  });
  int total;
  const int before = block_exclusive_scan<BAR, NTHR, TID0>(cnt_gt | (cnt_eq << 16), sm.warp_scan, total);
  const int need_eq = k - (total & 0xffff);  // ties to take
  const int eq_before = before >> 16;
  int pos = (before & 0xffff) + min(eq_before, need_eq);
  int eq_left = max(0, min(cnt_eq, need_eq - eq_before));
  for_each_word([&](int col, uint32_t w) {
    const uint32_t lo = w & 0xffffu, hi = w >> 16;
    bool take = lo > thresh;
    if (lo == thresh && lo && eq_left > 0) take = true, --eq_left;
    if (take) sm.sel[pos++] = col;
    take = hi > thresh;
    if (hi == thresh && hi && eq_left > 0) take = true, --eq_left;
    if (take) sm.sel[pos++] = col + 1;
  });
  // last page (quest.py:1888-1889: safe_num_pages - 1)
  const int num_pages = (ctx + kPage - 1) / kPage;
  if (tid == 0) sm.sel[k] = max(num_pages, 1) - 1;
  group_sync<BAR, NTHR>();
  for (int j = tid; j <= k; j += NTHR) {
    const int page = sm.sel[j];
    sm.sel_slot[j] = pt_row[page];
    if (sel_pages_out) sel_pages_out[j] = page;
  }
  group_sync<BAR, NTHR>();
  // packed token slots: (k+1)*16 int32 = (k+1)*4 int4
  int4* out4 = reinterpret_cast<int4*>(out_row);
  for (int idx = tid; idx < (k + 1) * 4; idx += NTHR) {
    const int base = sm.sel_slot[idx >> 2] * kPage + (idx & 3) * 4;
    out4[idx] = make_int4(base, base + 1, base + 2, base + 3);
  }
  if (tid == 0) *len_out = k * kPage + (ctx - (num_pages - 1) * kPage);
  group_sync<BAR, NTHR>();
  if (tdbg && tid == 0) {
    const unsigned long long tq3 = global_timer_ns();
    tdbg[0] += tq1 - tq0, tdbg[1] += tq2 - tq1, tdbg[2] += tq3 - tq2;
  }
}

template <int BAR, int NTHR, int TID0 = 0>
__device__ void quest_select_pack_row(const uint16_t* __restrict__ scores_row, int n_cols, int n_valid, int k,
                                      const int32_t* __restrict__ pt_row, int ctx,
                                      int32_t* __restrict__ out_row, int32_t* __restrict__ len_out,
                                      int32_t* __restrict__ sel_pages_out, SelectSmem& sm,
                                      unsigned long long* tdbg = nullptr) {
  // Short rows spread over all threads (1 or 2 vectors each instead of 4 in a quarter of the threads): a search round
  // costs every warp the instructions of its VPT vectors, whether they hold scores or padding.
  const int nvec = (n_cols + 7) >> 3;
  if (nvec <= NTHR)
    quest_select_pack_impl<BAR, NTHR, 1, TID0>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out,
                                               sel_pages_out, sm, tdbg);
  else if (nvec <= 2 * NTHR)
    quest_select_pack_impl<BAR, NTHR, 2, TID0>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out,
                                               sel_pages_out, sm, tdbg);
  else if (nvec <= kSelectVecPerThread * NTHR)
    quest_select_pack_impl<BAR, NTHR, kSelectVecPerThread, TID0>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row,
                                                                 len_out, sel_pages_out, sm, tdbg);
  else
    quest_select_pack_impl<BAR, NTHR, 0, TID0>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out,
                                               sel_pages_out, sm, tdbg);
}

__global__ void __launch_bounds__(256) quest_select_pack_kernel(
    const uint16_t* __restrict__ scores, int64_t score_stride, int n_cols,
    const int32_t* __restrict__ page_table, int64_t pt_stride, const int32_t* __restrict__ req_rows,
    const int32_t* __restrict__ context_lens, int k, int32_t* __restrict__ packed_slots,
    int64_t out_stride, int32_t* __restrict__ sparse_lens, int32_t* __restrict__ selected_pages,
    const int32_t* __restrict__ dense_table, int64_t dense_stride, int token_budget,
    int page_budget_base, int out_width) {
  __shared__ SelectSmem sm;
  pdl_wait();
  const int b = blockIdx.x;
  const int ctx = context_lens[b];
  const int64_t row = req_rows[b];
  int32_t* out_row = packed_slots + (int64_t)b * out_stride;
  if (dense_table != nullptr) {
    const int num_pages = (ctx + kPage - 1) / kPage;
    const int32_t* dense_row = dense_table + row * dense_stride;
    if (ctx <= token_budget || num_pages <= page_budget_base) {
      // dense row (quest.py:1867-1868, 1903-1911)
      for (int i = threadIdx.x; i < out_width; i += blockDim.x) out_row[i] = dense_row[i];
      if (threadIdx.x == 0) sparse_lens[b] = ctx;
      if (selected_pages)
        for (int j = threadIdx.x; j <= k; j += blockDim.x)
          selected_pages[(int64_t)b * (k + 1) + j] = -1;
      return;
    }
    for (int i = (k + 1) * kPage + threadIdx.x; i < out_width; i += blockDim.x)
      out_row[i] = dense_row[i];
  }
  const int n_valid = max((ctx + kPage - 1) / kPage, 1) - 1;
  quest_select_pack_row<0, 256>(scores + (int64_t)b * score_stride, n_cols, min(n_valid, n_cols), k,
                                page_table + row * pt_stride, ctx, out_row, sparse_lens + b,
                                selected_pages ? selected_pages + (int64_t)b * (k + 1) : nullptr, sm);
}

// =========================================================================
// Page scoring (+ fused top-k / packing).
//
// A "unit" is 16 consecutive logical pages of one batch row; only valid pages
// (p < ceil(ctx/16) - 1) form units.  The bounds of a unit are one contiguous bounds chunk
// (bounds_layout.cuh), streamed as hs ring stages of hh = Hkv/hs heads each.  Units are handed out
// dynamically (sched.cuh): a scheduler warp claims chunks of ~64 KB from per-row counters and posts
// descriptors; ONE loader warp resolves each unit's chunk id through the unit table (fetched
// 1-4 descriptors ahead: ~64 KB worth) and issues one bulk async copy (cp.async.bulk, SASS UBLKCP) per ring
// stage, 8-32 KB each, counted on the stage's mbarrier.  Consumer warps take units round-robin;
// per unit and KV head:
//   A = bounds rows (16 pages x 128, ldmatrix from the swizzled stage), B = q+ / q- (128 x 8 heads)
//   score = bf16(bf16(A_max.q+) + bf16(A_min.q-)), then max over all local heads.
// A CTA leaving a row adds the units it scored to the row's counter; the row's designated CTA runs
// the top-k selection and slot packing for it in the same launch (8 dedicated select warps).
// =========================================================================
constexpr int kScoreConsumerWarps = 6;
constexpr int kScoreConsumerThreads = kScoreConsumerWarps * 32;
constexpr int kScoreLoaderWarp = kScoreConsumerWarps;      // one loader warp ...
constexpr int kScoreSchedWarp = kScoreConsumerWarps + 1;   // ... and one scheduler warp
constexpr int kIdDepth = 4;             // most descriptors whose chunk ids are in flight ahead of the copies (the table read
                                        // is an L2 hit at best: ~1 us under load, i.e. 40 KB of streaming per SM)
constexpr int kScoreSelectWarps = 8;    // dedicated top-k warps: the stream never waits for a selection
constexpr int kScoreSelectThreads = kScoreSelectWarps * 32;
constexpr int kScoreThreads = kScoreConsumerThreads + 2 * 32 + kScoreSelectThreads;
constexpr int kScoreMaxSmem = 227 * 1024;
static_assert(kIdDepth < kSchedQueue, "the loader reads up to kIdDepth descriptors ahead of the one it issues");

struct ScoreParams {
  const __nv_bfloat16* q;
  int64_t q_stride_b;
  const __nv_bfloat16* bounds;  // bounds chunks of this layer
  int64_t num_chunks;
  const int32_t* unit_table;  // [rows][ut_stride]: bounds chunk of (row, logical page / 16)
  int64_t ut_stride;
  const int32_t* page_table;  // fused select only: logical page -> physical page (slot packing)
  int64_t pt_stride;
  const int32_t* req_rows;
  const int32_t* context_lens;
  uint16_t* scores;
  int64_t score_stride;
  int batch, hq, hkv, n_cols, stages;
  int hsplit;  // ring stages per unit (bounds_head_split)
  unsigned long long* dbg_times;  // experiments: per-CTA timestamps (svb200_debug_times), else null
  int groups;                     // row groups of the scheduler
  int chunk_units;                // units per claimed chunk
  int min_batch, share_div;       // guided claim size = max(min_batch, left / (share_div * CTAs per group))
  int no_math;                    // experiments: consumers release the stages without the contractions
  int evict_first;                // the bounds stream's copies are tagged evict-first in L2
  int poll_first_ns;              // one-launch layer: first sleep of a producer warp that finds its row not selected yet
  int look;                       // loader look-ahead in descriptors (1, 2 or 4)
  int tail_chunks;                // the last tail_chunks chunks of a row are claimed one at a time and posted unit by unit
  int32_t* sched;  // sched_scratch_ints(batch) zeroed ints
  // fused select only
  int k;
  int32_t* packed_slots;
  int64_t out_stride;
  int32_t* sparse_lens;
  int32_t* selected_pages;
  // fused store of the step's new K/V rows (store.key == null: none), one token per batch row
  StoreArgs store;
  int store_tokens;
  // attention phase of a one-launch layer (AHPC > 0)
  AttnParams attn;
};

struct ScoreSmemLayout {
  int ring_off, q_off, bar_off, valid_off, rows_off, sched_off, select_off, total;
};

__host__ __device__ inline int score_stage_bytes(int hkv, int hsplit) { return (int)bounds_part_bytes(hkv, hsplit); }

__host__ __device__ inline ScoreSmemLayout score_smem_layout(int stages, int hq, int hkv, int hsplit, int batch,
                                                             bool fused, int q_copies) {
  ScoreSmemLayout L;
  int off = 0;
  L.ring_off = off;   off += stages * score_stage_bytes(hkv, hsplit);
  L.q_off = off;      off += q_copies * hq * kRowBytes;
  L.bar_off = off;    off += 2 * stages * 8;
  L.valid_off = off;  off += ((stages * 4 + 15) / 16) * 16;
  L.rows_off = off;   off += (2 * batch * 4 + 15) / 16 * 16;  // valid pages, table row of each batch row
  L.sched_off = off;  off += (int)((sizeof(SchedSmem) + 15) / 16 * 16);
  L.select_off = off; off += fused ? (int)((sizeof(SelectSmem) + 15) / 16 * 16) : 0;
  L.total = off;
  return L;
}

// bf16x2 clamp_min(0) / clamp_max(0) by sign bit (NaN keeps its sign's side, like the reference's clamp)
__device__ __forceinline__ uint32_t q_neg_mask(uint32_t x) { return ((x >> 15) & 0x00010001u) * 0xffffu; }
__device__ __forceinline__ uint32_t q_pos(uint32_t x) { return x & ~q_neg_mask(x); }
__device__ __forceinline__ uint32_t q_neg(uint32_t x) { return x & q_neg_mask(x); }

__device__ __forceinline__ void score_consumer_sync() { group_sync<1, kScoreConsumerThreads>(); }

// The select group's selection of one row, as a function of its own (one-launch layer kernels).
// (its scratch is named by its offset in the kernel's dynamic shared memory, not by a pointer: through a generic
// pointer the compiler could neither use shared-memory instructions nor reorder around them, and the selection took
// twice as long)
__device__ __noinline__ void fused_select_pack_row(const uint16_t* scores_row, int n_cols, int n_valid, int k,
                                                   const int32_t* pt_row, int ctx, int32_t* out_row, int32_t* len_out,
                                                   int32_t* sel_pages_out, int select_off, unsigned long long* tdbg) {
  extern __shared__ __align__(128) unsigned char smem[];
  quest_select_pack_row<2, kScoreSelectThreads, (kScoreThreads - kScoreSelectThreads)>(
      scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out, sel_pages_out,
      *reinterpret_cast<SelectSmem*>(smem + select_off), tdbg);
}

// Attention phase of a one-launch layer (AHPC > 0 in quest_page_score_kernel, whose header describes it); entered by
// warps 0-11 of a CTA.  dbg (experiments): slots 16.. of the CTA's time stamps.
template <int AHPC>
__device__ __forceinline__ void score_attention_phase(const ScoreParams& p, unsigned char* smem, int* s_flag,
                                                      unsigned long long* dbg) {
  const AttnParams& ap = p.attn;
  const int B = p.batch;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int* sel_done = p.sched + 1 + 2 * B;
  const uint32_t tuning_poll_ns = p.poll_first_ns;
  auto wait_row = [&](int b) {  // whole warp; one lane polls
    if ((tid & 31) == 0) {
      SpinGuard guard;
      // few looks, far apart: a selection takes >= 6 us from the moment the last CTA leaves the row, and hundreds of
      // warps polling the flags' cache lines (shared with the rows' counters) slowed the running selections down
      uint32_t ns = tuning_poll_ns;
      while (ld_acquire_gpu(sel_done + b) == 0) {
        __nanosleep(ns);
        ns = 1000;
        guard.tick(kSpinLimitNs);
      }
      if (dbg && warp == kConsumerWarps && dbg[23] == 0) dbg[23] = global_timer_ns();
    }
    __syncwarp();
  };
  // tokens of batch entry b the attention reads: what the selection writes to sparse_lens (known up front)
  auto len_of = [&](int b) {
    const int ctx = p.context_lens[b];
    const int pages = (ctx + kPage - 1) / kPage;
    return ctx > 0 ? min(p.k * kPage + ctx - (pages - 1) * kPage, ap.max_len) : 0;
  };
  if (warp <= kScoreSchedWarp) {
    // this CTA's scoring is over: its consumers drained every stage, its loader and scheduler are done
    group_sync<3, kConsumerThreads>();
    if (dbg && tid == 0) dbg[16] = global_timer_ns();
    attn_setup<AHPC, 1>(ap, smem, len_of);
    __threadfence_block();
    asm volatile("bar.arrive 4, %0;" ::"n"(kConsumerThreads + kProducerWarps * 32) : "memory");
    // ... and wait, blocked in hardware, until the producers are in (in a selecting CTA they are select warps
    // until their rows are done): the consumers do not poll their ring barriers next to a running selection
    asm volatile("bar.sync 5, %0;" ::"n"(kConsumerThreads + kProducerWarps * 32) : "memory");
    if (dbg && tid == 0) dbg[17] = global_timer_ns();
  } else {
    asm volatile("bar.sync 4, %0;" ::"n"(kConsumerThreads + kProducerWarps * 32) : "memory");
    asm volatile("bar.arrive 5, %0;" ::"n"(kConsumerThreads + kProducerWarps * 32) : "memory");
  }
  // one call site for both roles: the kernel holds ONE copy of the attention code (two inlined copies made it 480 KB)
  attn_main<AHPC, 1, false, 3, true>(ap, smem, wait_row, dbg ? dbg + 16 : nullptr);
  if (tid == 0 && atomicAdd(s_flag, 1) == 1) sched_finish(p.sched, B);
}

// QSPLIT: q+ and q- are both staged (when that does not cost a ring stage); else q is staged once and
// split by sign when a fragment is used.
// AHPC > 0 (with FUSED_SELECT): the launch goes on with the layer's sparse decode attention (decode_attn.cuh, AHPC
// KV heads per CTA) -- one launch per layer.  A CTA whose scoring work is done re-uses its ring + q area for the
// attention ring: warps 0-7 (consumers, loader, scheduler) become the attention consumers and do its set-up
// while the selections of the last rows are still running; warps 8-11 become the attention producers once their
// own selections are done, and wait, per batch entry, for the entry's selection flag (set with release semantics
// by the entry's selecting CTA) before they read its packed slots.  What is overlapped: the launch gap, the
// attention's set-up and its ramp with the selection tail of the scoring launch.  Nobody waits before its own
// scoring / selecting work is done and all CTAs are co-resident (grid <= SMs, one CTA per SM), so the flags
// cannot dead-lock.
template <int NT, bool FUSED_SELECT, bool QSPLIT, int AHPC = 0>
__global__ void __launch_bounds__(kScoreThreads, 1) quest_page_score_kernel(const __grid_constant__ ScoreParams p) {
  static_assert(AHPC == 0 || (FUSED_SELECT && NT == 1), "the attention phase follows a fused selection, group size <= 8");
  extern __shared__ __align__(128) unsigned char smem[];
  const int S = p.stages, hq = p.hq, hkv = p.hkv, B = p.batch;
  const int gsz = hq / hkv;
  const int hs = p.hsplit, hh = hkv / hs;     // stages per unit, KV heads per stage
  const int rec_stride = hh * 2 * kRowBytes;  // one page inside a stage: max rows of the stage's heads, then min rows
  const int stage_bytes = kPage * rec_stride;
  const ScoreSmemLayout L = score_smem_layout(S, hq, hkv, hs, B, FUSED_SELECT, QSPLIT ? 2 : 1);
  unsigned char* ring = smem + L.ring_off;
  // sq[sign][head][kk][t] : uint2 {b0, b1} B fragments of q+ / q-, kk XOR-swizzled by head (QSPLIT), or
  // sq[head][kk][t] of q itself, split by sign when a fragment is used (two copies cost a ring stage at 64 heads)
  uint32_t* sq = reinterpret_cast<uint32_t*>(smem + L.q_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* empty_bar = full_bar + S;
  uint32_t* sValid = reinterpret_cast<uint32_t*>(smem + L.valid_off);
  int* sNv = reinterpret_cast<int*>(smem + L.rows_off);  // valid previous pages of row b
  int* sReq = sNv + B;                                   // table row of batch row b
  SchedSmem& sched = *reinterpret_cast<SchedSmem*>(smem + L.sched_off);
  __shared__ int s_flag;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long* dbg = p.dbg_times ? p.dbg_times + (size_t)blockIdx.x * (AHPC > 0 ? 32 : 16) : nullptr;

  // ---- setup that touches no global memory (runs while a programmatic predecessor drains)
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);  // the loader's expect_tx arrival; the copy's bytes complete the phase
      mbar_init(&empty_bar[s], 1);
      // no fill yet: a consumer whose first unit is a later round of the stage passes the fresh barrier's
      // parity test and must not find a tag a previous launch left in shared memory (tags are 15 bits)
      sValid[s] = 0xffffffffu;
    }
    sched_init(sched, 1 + kScoreConsumerWarps);
    s_flag = 0;
    mbar_fence_init();
  }
  pdl_wait();
  if (dbg && tid == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    dbg[0] = global_timer_ns();
    dbg[7] = smid;
    dbg[2] = dbg[4] = dbg[5] = dbg[6] = dbg[13] = dbg[14] = dbg[15] = 0;
    if constexpr (AHPC > 0)
      for (int x = 16; x < 32; ++x) dbg[x] = 0;
  }
  for (int b = tid; b < B; b += kScoreThreads) {
    const int ctx = p.context_lens[b];
    sNv[b] = min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);
    sReq[b] = p.req_rows[b];
  }
  // columns no unit covers are -inf (quest.py:1883-1885)
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const int ctx = p.context_lens[b];
    const int nv = min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);
    uint16_t* srow = p.scores + (int64_t)b * p.score_stride;
    for (int c = (nv + kPage - 1) / kPage * kPage + tid; c < p.n_cols; c += kScoreThreads) srow[c] = 0xff80u;
  }
  __syncthreads();
  pdl_launch_dependents();
  if (dbg && tid == 0) dbg[1] = global_timer_ns();
  auto units_of_row = [&](int r) { return (sNv[r] + kPage - 1) / kPage; };
  int* next_chunk = p.sched + 1;
  int* units_done = p.sched + 1 + B;

  if (warp == kScoreSchedWarp) {
    // ============================================================== scheduler
    const int depth = min(kSchedQueue, (S / hs + p.chunk_units - 1) / p.chunk_units + p.look + 1);
    sched_run(sched, next_chunk, B, p.groups, p.chunk_units, p.min_batch, p.share_div, depth, units_of_row,
              ReserveNone(), (dbg && blockIdx.x >= (unsigned)B) ? dbg + 15 : nullptr, p.tail_chunks);
    if constexpr (AHPC == 0) return;
  }

  if (warp == kScoreLoaderWarp) {
    // ============================================================== loader
    // Stage-load ii = i * hs + part (i: CTA-local running unit index) goes into stage ii % S.  Lane l holds
    // the bounds chunk of unit unit_begin + l of a descriptor, fetched LOOK descriptors ahead, and issues ONE bulk
    // async copy per stage-load of its unit.  A partly valid last unit is copied whole (the chunk
    // exists; rows of pages that do not count are masked when the scores are written).
    const uint32_t part_bytes = (uint32_t)stage_bytes;
    unsigned long long w_desc = 0, w_empty = 0;
    // LOOK = descriptors whose chunk ids are in flight ahead of the copies: enough to cover a table read (~1 us, i.e.
    // ~64 KB of streaming per SM) and no more, because what is posted ahead is committed to this CTA
    auto run_loader = [&](auto look_tag) {
    constexpr int LOOK = decltype(look_tag)::value;
    ChunkDesc dq[LOOK];
    int idq[LOOK];
    int n_read = 0;
    bool ended = false;
    auto fetch = [&](ChunkDesc& d, int& id) {
      id = -1;
      if (ended) {
        d.row = -1;
        return;
      }
      const unsigned long long tw = dbg ? global_timer_ns() : 0;
      d = sched_read(sched, n_read++);
      if (dbg) w_desc += global_timer_ns() - tw;
      if (d.row < 0) {
        ended = true;
        return;
      }
      if (lane < d.units()) id = __ldg(p.unit_table + (int64_t)sReq[d.row] * p.ut_stride + d.unit_begin + lane);
    };
#pragma unroll
    for (int t = 0; t < LOOK; ++t) fetch(dq[t], idq[t]);
    const char* bounds_bytes = reinterpret_cast<const char*>(p.bounds);
    const uint64_t l2_policy = l2_evict_first_policy();
    // dq / idq form a circular FIFO indexed at compile time (the loop body is unrolled LOOK
    // times) and a slot is refilled only AFTER its id was used: the register that waits for a table
    // load is then never copied or renamed, so the load's latency is only paid when the id is finally
    // used, LOOK descriptors later.  (Refilling the slot before the use made the compiler keep the
    // old value and move the new one into place at the loop's back edge: a wait for the load that was
    // issued last, i.e. one full table-read latency per descriptor, which capped launches with small
    // units (<= 16 KB, TP shards) at ~2.3 units/us per SM whatever their size.)
    bool done = false;
    while (!done) {
#pragma unroll
      for (int u = 0; u < LOOK; ++u) {
        if (done) break;
        const ChunkDesc cur = dq[u];
        const int ids = idq[u];
        if (cur.row < 0) {
          done = true;
          break;
        }
        // Lane j feeds unit j of the descriptor (its own table entry, its own stages; a descriptor never holds
        // more units than the ring has room for, so the lanes' stages are distinct): the instructions of a feed
        // -- stage arithmetic, barrier poll, tag, expect_tx, copy -- are paid once per descriptor, not once per
        // unit.  One lane feeding unit after unit took ~0.4 us per unit, which capped launches with 8-16 KB
        // units (1-2 KV heads per rank) at 2.3 units/us per SM whatever their size.
        if (lane < cur.units()) {
          const int c = ids;
          const bool have = c >= 0 && c < p.num_chunks;
          const int cnt = sNv[cur.row] - (cur.unit_begin + lane) * kPage;  // valid pages of the unit (> 0)
          const uint32_t umask = have ? (cnt >= kPage ? 0xffffu : ((1u << max(cnt, 0)) - 1u)) : 0u;
          for (int part = 0; part < hs; ++part) {
            const int ii = (cur.unit_base + lane) * hs + part;
            const int s = ii % S;
            const uint32_t round = (uint32_t)(ii / S);
            const unsigned long long te = dbg ? global_timer_ns() : 0;
            mbar_wait(&empty_bar[s], (round & 1u) ^ 1u);
            if (dbg) w_empty += global_timer_ns() - te;
            // valid pages of the unit | the stage's fill count (the consumer checks it: see there)
            sValid[s] = umask | ((round & 0x7fffu) << 16);
            if (have) {
              mbar_arrive_expect_tx(&full_bar[s], part_bytes);
              if (p.evict_first)
                bulk_g2s_hint(smem_u32(ring + s * stage_bytes),
                              bounds_bytes + ((int64_t)c * hs + part) * (int64_t)part_bytes, part_bytes, &full_bar[s], l2_policy);
              else
              bulk_g2s(smem_u32(ring + s * stage_bytes),
                       bounds_bytes + ((int64_t)c * hs + part) * (int64_t)part_bytes, part_bytes, &full_bar[s]);
            } else {
              mbar_arrive(&full_bar[s]);
            }
          }
        }
        __syncwarp();
        fetch(dq[u], idq[u]);
      }
    }
    };  // run_loader
    if (p.look >= 4) run_loader(std::integral_constant<int, 4>());
    else if (p.look >= 2) run_loader(std::integral_constant<int, 2>());
    else run_loader(std::integral_constant<int, 1>());
    if (dbg && lane == 0) dbg[4] = global_timer_ns(), dbg[8] = w_desc, dbg[9] = w_empty;
    if constexpr (AHPC == 0) return;
  }

  if (warp > kScoreSchedWarp) {
    // ============================================================== select warps
    // First, while the stream starts: pull the unit table of every row into L2 (each 128-byte line by one thread
    // of the launch).  The loaders read it a few descriptors ahead of their copies; from DRAM that look-ahead
    // was too short for small units (one KV head per rank: 8 KB per unit, 2.3 units/us per SM).
    {
      const int gtid = blockIdx.x * kScoreSelectThreads + (tid - (kScoreThreads - kScoreSelectThreads));
      const int gthreads = gridDim.x * kScoreSelectThreads;
      for (int b = 0; b < B; ++b) {
        const int lines = (units_of_row(b) * 4 + 127) >> 7;
        const char* row = reinterpret_cast<const char*>(p.unit_table + (int64_t)sReq[b] * p.ut_stride);
        for (int l = (gtid + b * 61) % gthreads; l < lines; l += gthreads)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(row + (int64_t)l * 128));
      }
    }
    // CTA c selects rows c, c + gridDim, ...: it waits until the row's counter shows every unit
    // scored (the consumers of all CTAs add what they scored when they leave a row), then runs the
    // top-k and packing.  The stream (loader / consumers) never waits for a selection.
    if constexpr (FUSED_SELECT) {
      SelectSmem* sel_sm = reinterpret_cast<SelectSmem*>(smem + L.select_off);
      const int stid = tid - (kScoreThreads - kScoreSelectThreads);
      // The step's new K/V rows (and the bounds of a page they complete) are written here, by warps
      // that have nothing else to do until the first row is scored.  The token's page is the row's
      // last one, which this launch never scores (n_valid = pages - 1); attention runs after the launch.
      if (p.store.key != nullptr)
        for (int tok = blockIdx.x; tok < p.store_tokens; tok += gridDim.x)
          store_kv_token(p.store, tok, hkv, stid, kScoreSelectThreads);
      for (int b = blockIdx.x; b < B; b += gridDim.x) {
        const int need = units_of_row(b);
        // while the row is being scored: pull its page-table row into L2, so that the look-ups of the selected
        // pages at the very end of the launch do not pay a DRAM round trip
        {
          const char* pt_bytes = reinterpret_cast<const char*>(p.page_table + (int64_t)sReq[b] * p.pt_stride);
          for (int off = stid * 128; off < (sNv[b] + 1) * 4; off += kScoreSelectThreads * 128)
            asm volatile("prefetch.global.L2 [%0];" ::"l"(pt_bytes + off));
        }
        if (need > 0 && stid == 0) {
          const volatile int* ctr = units_done + b;
          SpinGuard guard;
          while (*ctr < need) {
            __nanosleep(64);
            guard.tick(kSpinLimitNs);
          }
        }
        group_sync<2, kScoreSelectThreads>();
        __threadfence();
        const unsigned long long ts = dbg ? global_timer_ns() : 0;
        if constexpr (AHPC > 0)  // (its own register allocation: inlined next to the attention phase it spilled)
          fused_select_pack_row(p.scores + (int64_t)b * p.score_stride, p.n_cols, sNv[b], p.k,
                                p.page_table + (int64_t)sReq[b] * p.pt_stride, p.context_lens[b],
                                p.packed_slots + (int64_t)b * p.out_stride, p.sparse_lens + b,
                                p.selected_pages ? p.selected_pages + (int64_t)b * (p.k + 1) : nullptr, L.select_off,
                                dbg ? dbg + 13 : nullptr);
        else
        quest_select_pack_row<2, kScoreSelectThreads, (kScoreThreads - kScoreSelectThreads)>(
            p.scores + (int64_t)b * p.score_stride, p.n_cols, sNv[b], p.k,
            p.page_table + (int64_t)sReq[b] * p.pt_stride, p.context_lens[b],
            p.packed_slots + (int64_t)b * p.out_stride, p.sparse_lens + b,
            p.selected_pages ? p.selected_pages + (int64_t)b * (p.k + 1) : nullptr, *sel_sm,
            dbg ? dbg + 13 : nullptr);
        if (dbg && stid == 0) dbg[5] += global_timer_ns() - ts;
        if constexpr (AHPC > 0) {
          // the row's packed slots (and this CTA's earlier K/V store) are published: the select group's last barrier
          // ordered every thread's writes before this thread's fence
          if (stid == 0) {
            __threadfence();
            st_release_gpu(p.sched + 1 + 2 * B + b, 1);
          }
        }
      }
      // the CTA's scratch may be recycled once BOTH its consumers and its select warps are done
      if (stid == 0 && atomicAdd(&s_flag, 1) == 1) sched_finish(p.sched, B);
    }
    if constexpr (AHPC == 0) return;
  }

  if (warp < kScoreConsumerWarps) {
  // ================================================================ consumers
  const int g = lane >> 2, t = lane & 3;
  const uint2* sq2 = reinterpret_cast<const uint2*>(sq);
  const int neg_off = hq * 32;  // uint2 offset of the q- fragments (QSPLIT)
  // q of row b in B-fragment order (q+ / q- = clamp_min(0) / clamp_max(0), quest.py:1796-1797, see q_pos / q_neg)
  auto stage_q = [&](int b) {
    const uint4* qrow = reinterpret_cast<const uint4*>(p.q + (int64_t)b * p.q_stride_b);
    for (int c = tid; c < hq * 16; c += kScoreConsumerThreads) {
      const int head = c >> 4, chunk = c & 15;
      const uint4 x = qrow[c];
      const uint32_t w[4] = {x.x, x.y, x.z, x.w};
      const int kk = chunk >> 1, hf = chunk & 1;
      const int base = ((head * 8 + (kk ^ (head & 7))) * 4) * 2 + hf;
#pragma unroll
      for (int tt = 0; tt < 4; ++tt) {
        if constexpr (QSPLIT) {
          sq[base + tt * 2] = q_pos(w[tt]);
          sq[neg_off * 2 + base + tt * 2] = q_neg(w[tt]);
        } else {
          sq[base + tt * 2] = w[tt];
        }
      }
    }
  };

  // Unit i goes to warp i % NW, which drains all of its hs stage-loads.  A parity wait cannot tell a
  // phase from the one two before it, and a stage is drained by different warps in consecutive
  // rounds (NW need not divide S / hs): the loader therefore tags the stage's valid mask with its
  // fill count, and a warp only trusts the barrier once the tag shows the fill it is waiting for.
  constexpr int NW = kScoreConsumerWarps;
  // ldmatrix lane -> (page row, 16-byte column half); the stage is swizzled by (page & 7) (bounds_layout.cuh)
  const uint32_t lm_row = (uint32_t)((lane & 7) + ((lane >> 3) & 1) * 8) * (uint32_t)rec_stride;
  const uint32_t lm_hi = (uint32_t)(lane >> 4) << 4, lm_xor = (uint32_t)(lane & 7) << 4;
  bool first_seen = false;
  int cur_row = -2, visit_units = 0, total_units = 0;  // -2: no row yet (-1 is the end descriptor)
  unsigned long long c_desc = 0, c_full = 0;
  for (int n = 0;; ++n) {
    const unsigned long long tc = dbg ? global_timer_ns() : 0;
    const ChunkDesc d = sched_read(sched, n);
    if (dbg) c_desc += global_timer_ns() - tc;
    if (d.row != cur_row) {
      // leave cur_row: add the units scored here to the row's counter (fire and forget; the row's
      // selecting CTA polls it), stage the next row's q
      score_consumer_sync();
      if constexpr (FUSED_SELECT) {
        if (tid == 0 && cur_row >= 0) {
          __threadfence();
          atomicAdd(units_done + cur_row, visit_units);
        }
      }
      if (d.row >= 0) stage_q(d.row);
      score_consumer_sync();
      cur_row = d.row;
      visit_units = 0;
      if (d.row < 0) break;
    }
    visit_units += d.units();
    total_units += d.units();
    uint16_t* srow = p.scores + (int64_t)d.row * p.score_stride;
    for (int j = 0; j < d.units(); ++j) {
      const int i = d.unit_base + j;
      if (i % NW != warp) continue;
      float best0 = -INFINITY, best1 = -INFINITY;
      uint32_t mask = 0;
      for (int part = 0; part < hs; ++part) {
        const int ii = i * hs + part;
        const int s = ii % S;
        const uint32_t round = (uint32_t)(ii / S);
        const unsigned long long tf = dbg ? global_timer_ns() : 0;
        while (true) {
          mbar_wait(&full_bar[s], round & 1u);
          mask = (uint32_t)ld_volatile_shared(reinterpret_cast<const int*>(sValid + s));
          if ((mask >> 16) == (round & 0x7fffu)) {
            mbar_wait(&full_bar[s], round & 1u);  // the tag is written before the copy lands: wait for it
            break;
          }
          __nanosleep(64);
        }
        if (dbg) c_full += global_timer_ns() - tf;
        if (dbg && tid == 0 && !first_seen) dbg[2] = global_timer_ns(), first_seen = true;
        const uint32_t abase = smem_u32(ring + s * stage_bytes) + lm_row;
        // KV heads two at a time: the two heads' ldmatrix -> mma chains are independent, so the in-order
        // issue of one warp overlaps their latencies (one head at a time left the consumers the bottleneck
        // at Hkv 8); an odd last head (and the only head of a TP-8 shard) runs alone
        const int hbase = part * hh;  // first KV head of this stage
        auto score_heads = [&](auto nh_tag, int h) {
          constexpr int NH = decltype(nh_tag)::value;
          float cp[NH][NT][4], cn[NH][NT][4];
#pragma unroll
          for (int x = 0; x < NH; ++x)
#pragma unroll
            for (int nt = 0; nt < NT; ++nt)
#pragma unroll
              for (int e = 0; e < 4; ++e) cp[x][nt][e] = cn[x][nt][e] = 0.f;
          uint32_t amax[NH], amin[NH];
#pragma unroll
          for (int x = 0; x < NH; ++x) amax[x] = abase + (h + x) * kRowBytes, amin[x] = amax[x] + hh * kRowBytes;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            const uint32_t koff = ((uint32_t)(kk * 32) | lm_hi) ^ lm_xor;
            uint32_t ax[NH][4], an[NH][4];
#pragma unroll
            for (int x = 0; x < NH; ++x) {
              ldmatrix_x4(ax[x], amax[x] + koff);
              ldmatrix_x4(an[x], amin[x] + koff);
            }
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
#pragma unroll
              for (int x = 0; x < NH; ++x) {
                const int qh = (hbase + h + x) * gsz + min(nt * 8 + g, gsz - 1);
                const int f = (qh * 8 + (kk ^ (qh & 7))) * 4 + t;
                uint2 bp = sq2[f], bn;
                if constexpr (QSPLIT) {
                  bn = sq2[neg_off + f];
                } else {
                  bn = make_uint2(q_neg(bp.x), q_neg(bp.y));
                  bp = make_uint2(q_pos(bp.x), q_pos(bp.y));
                }
                mma_16816(cp[x][nt], ax[x], bp.x, bp.y);
                mma_16816(cn[x][nt], an[x], bn.x, bn.y);
              }
            }
          }
          // c0:(page g, head 2t) c1:(g, 2t+1) c2:(g+8, 2t) c3:(g+8, 2t+1)
#pragma unroll
          for (int x = 0; x < NH; ++x) {
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
              float sc[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) sc[e] = round_bf16(round_bf16(cp[x][nt][e]) + round_bf16(cn[x][nt][e]));
              const int col0 = nt * 8 + 2 * t;
              if (col0 < gsz) best0 = fmaxf(best0, sc[0]), best1 = fmaxf(best1, sc[2]);
              if (col0 + 1 < gsz) best0 = fmaxf(best0, sc[1]), best1 = fmaxf(best1, sc[3]);
            }
          }
        };
        if (!p.no_math) {
          int h = 0;
          for (; h + 1 < hh; h += 2) score_heads(std::integral_constant<int, 2>(), h);
          if (h < hh) score_heads(std::integral_constant<int, 1>(), h);
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[s]);
      }  // part
      best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 1));
      best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 2));
      best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 1));
      best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 2));
      if (t == 0) {
        const int p0 = (d.unit_begin + j) * kPage + g, p1 = p0 + 8;
        if (p0 < p.n_cols) {
          __nv_bfloat16 v = __float2bfloat16_rn(best0);
          srow[p0] = ((mask >> g) & 1u) ? *reinterpret_cast<uint16_t*>(&v) : (uint16_t)0xff80u;
        }
        if (p1 < p.n_cols) {
          __nv_bfloat16 v = __float2bfloat16_rn(best1);
          srow[p1] = ((mask >> (g + 8)) & 1u) ? *reinterpret_cast<uint16_t*>(&v) : (uint16_t)0xff80u;
        }
      }
    }
  }
  if (tid == 0) {
    if (dbg) dbg[3] = global_timer_ns(), dbg[6] = (unsigned long long)total_units, dbg[11] = c_desc, dbg[12] = c_full;
    if constexpr (AHPC == 0)
      if (!FUSED_SELECT || atomicAdd(&s_flag, 1) == 1) sched_finish(p.sched, B);
  }
  }  // consumers

  if constexpr (AHPC > 0) {
    if (warp <= kScoreSchedWarp + kProducerWarps) score_attention_phase<AHPC>(p, smem, &s_flag, dbg);
  }
}

// =========================================================================
// Host launchers
// =========================================================================
int launch_quest_page_minmax(const void* k_cache, void* bounds_layer, const int32_t* page_rec,
                             const int32_t* slot_mapping, int n_tokens, int hkv, int64_t num_slots,
                             int64_t num_chunks, cudaStream_t stream) {
  if (n_tokens <= 0) return 0;
  const int warps_per_block = 8;
  int blocks = (n_tokens + warps_per_block - 1) / warps_per_block;
  if (blocks > 148 * 8) blocks = 148 * 8;
  const cudaError_t e = launch_kernel(quest_page_minmax_kernel, dim3(blocks), dim3(warps_per_block * 32), 0, stream,
                                      (const __nv_bfloat16*)k_cache, (__nv_bfloat16*)bounds_layer, page_rec,
                                      slot_mapping, n_tokens, hkv, num_slots, num_chunks * kPage);
  return e != cudaSuccess ? cuda_error("quest_page_minmax", e) : check_launch("quest_page_minmax");
}

int launch_quest_decode_page_complete(const void* k_cache_all, void* bounds_all, const int32_t* page_rec,
                                      const int32_t* page_table, int64_t pt_stride,
                                      const int32_t* req_rows, const int32_t* seq_lens, int batch,
                                      int layers, int hkv, int64_t num_slots, int64_t num_pages,
                                      int64_t num_chunks, cudaStream_t stream) {
  if (batch <= 0 || layers <= 0) return 0;
  const cudaError_t e = launch_kernel(quest_decode_page_complete_kernel, dim3(batch, layers), dim3(128), 0, stream,
                                      (const __nv_bfloat16*)k_cache_all, (__nv_bfloat16*)bounds_all, page_rec,
                                      page_table, pt_stride, req_rows, seq_lens, hkv, num_slots, num_pages,
                                      num_chunks * kPage);
  return e != cudaSuccess ? cuda_error("quest_decode_page_complete", e) : check_launch("quest_decode_page_complete");
}

int launch_quest_select_pack(const void* scores, int64_t score_stride, int n_cols,
                             const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
                             const int32_t* context_lens, int batch, int k, int32_t* packed_slots,
                             int64_t out_stride, int32_t* sparse_lens, int32_t* selected_pages,
                             const int32_t* dense_table, int64_t dense_stride, int token_budget,
                             int page_budget_base, int out_width, cudaStream_t stream) {
  if (batch <= 0) return 0;
  const cudaError_t e = launch_kernel(quest_select_pack_kernel, dim3(batch), dim3(256), 0, stream,
                                      (const uint16_t*)scores, score_stride, n_cols, page_table, pt_stride, req_rows,
                                      context_lens, k, packed_slots, out_stride, sparse_lens, selected_pages,
                                      dense_table, dense_stride, token_budget, page_budget_base, out_width);
  return e != cudaSuccess ? cuda_error("quest_select_pack", e) : check_launch("quest_select_pack");
}

// Ring stages next to `q_copies` staged copies of q (0: no room).
static int score_stages(const ScoreParams& p, bool fused, int q_copies) {
  const int fixed = score_smem_layout(0, p.hq, p.hkv, p.hsplit, p.batch, fused, q_copies).total + 256;  // + static smem
  int stages = (kScoreMaxSmem - fixed) / (score_stage_bytes(p.hkv, p.hsplit) + 20);
  if (stages > 24) stages = 24;
  return stages < 2 ? 0 : stages;
}

template <int NT, bool FUSED, bool QSPLIT, int AHPC>
static cudaError_t launch_score_kernel(const ScoreParams& p, int grid, size_t smem, cudaStream_t stream) {
  auto kern = quest_page_score_kernel<NT, FUSED, QSPLIT, AHPC>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return e;
  return launch_kernel(kern, dim3(grid), dim3(kScoreThreads), smem, stream, p);
}

template <int NT, bool FUSED, int AHPC = 0>
static int launch_score_variant(ScoreParams p, cudaStream_t stream) {
  p.hsplit = bounds_head_split(p.hkv);
  const int stage_bytes = score_stage_bytes(p.hkv, p.hsplit);
  const int one = score_stages(p, FUSED, 1), two = score_stages(p, FUSED, 2);
  if (one == 0) {
    set_error("quest_page_score: %d q heads / %d kv heads / batch %d leave no shared memory for the bounds ring",
              p.hq, p.hkv, p.batch);
    return 1;
  }
  // q+ and q- are both staged only if that costs nothing of the ring
  bool qsplit = two == one;
  if (tuning(kTuneScoreQ) == 1) qsplit = false;
  if (tuning(kTuneScoreQ) == 2 && two > 0) qsplit = true;
  p.stages = qsplit ? two : one;
  // Row groups (sched.cuh): by default every row is its own group, i.e. gridDim/batch CTAs start on
  // each row and q is staged once per CTA; fewer groups make rows end at staggered times.
  int groups = tuning(kTuneScoreGroups);
  if (groups <= 0) groups = p.batch;
  if (groups > p.batch) groups = p.batch;
  if (groups < 1) groups = 1;
  p.groups = groups;
  const int64_t units_row = (p.n_cols + kPage - 1) / kPage;
  int grid = max_ctas();
  // What is claimed cannot be stolen, so claims bound the raggedness of the launch's tail; but a claim's round
  // trip costs 3-4 us under load and must hide behind queued streaming.  A claim is worth 1/8 of a CTA's share
  // of the launch, between 32 KB (small launches: TP shards, one-row batches) and 128 KB, made of two chunks.
  {
    const int64_t unit_bytes = (int64_t)stage_bytes * p.hsplit;
    const int64_t per_cta = (int64_t)p.batch * units_row * unit_bytes / grid;
    int64_t claim = per_cta / 8;
    claim = claim < 32 * 1024 ? 32 * 1024 : (claim > 128 * 1024 ? 128 * 1024 : claim);
    // A chunk is one descriptor: the scheduler posts it, every consumer warp reads it and the loader's lanes feed
    // its units in parallel; that handling costs ~0.4 us per descriptor whatever it carries (one unit per
    // descriptor: 54.7 us at the TP-8 shard shape, eight: 32.2 us; at cfg5 one 64 KB unit: 98.9 us, three: 95.7 us).
    // So a chunk carries 64 KB of small units (<= 16 KB: 1-2 KV heads per rank) and 192 KB of large ones.
    const int64_t chunk_target = unit_bytes < 32 * 1024 ? 64 * 1024 : 192 * 1024;
    int cu = tuning(kTuneChunkUnits) > 0 ? tuning(kTuneChunkUnits) : (int)((chunk_target + unit_bytes - 1) / unit_bytes);
    const int ring_units = p.stages / p.hsplit;  // the units of one descriptor occupy distinct ring stages
    if (cu > ring_units) cu = ring_units;
    p.chunk_units = cu < 1 ? 1 : (cu > kMaxChunkUnits ? kMaxChunkUnits : cu);
    const int64_t chunk_bytes = p.chunk_units * unit_bytes;
    int mb = (int)((claim + chunk_bytes - 1) / chunk_bytes);
    // (at least two chunks per claim while a chunk is short against the claim's round trip; a 128+ KB chunk is not:
    // cfg5 93.0 -> 92.0 us, cfg3 95.7 -> 93.4 with single-chunk claims)
    if (mb < kMinBatch && chunk_bytes < 128 * 1024) mb = kMinBatch;
    if (mb < 1) mb = 1;
    p.min_batch = tuning(kTuneMinBatch) > 0 ? tuning(kTuneMinBatch) : mb;
    // look-ahead of the loader: ~64 KB of descriptors (a table read is ~1 us), rounded to 1 / 2 / 4
    const int64_t want_look = (64 * 1024 + chunk_bytes - 1) / chunk_bytes;
    p.look = tuning(kTuneLook) > 0 ? tuning(kTuneLook) : (want_look >= 3 ? 4 : (want_look >= 2 ? 2 : 1));
    // Tail of a row (knob 16 = chunks + 1): what a CTA holds (ring + descriptors posted ahead + its claim) cannot be
    // stolen, and the CTAs' finishing times are 14 us apart at cfg5.  Claiming the last chunks of a row one at a time
    // and posting them one unit per descriptor was measured and is OFF: the extra descriptors cost more than the
    // finer tail saves (cfg5 95.1 vs 94.0 us, cfg3 98.7 vs 94.8, N=2 shard 58.5 vs 55.2 with 2 chunks per sharing CTA).
    p.tail_chunks = tuning(kTuneTailChunks) > 0 ? tuning(kTuneTailChunks) - 1 : 0;
  }
  // Static first piece = 1/share_div of a CTA's fair share of its home row.  A launch that gives a CTA less than
  // ~1 MB (TP shards, short batches) is over before guided claims could rebalance anything (what a CTA holds in
  // its ring and descriptor queue is most of its share): its rows are cut statically and evenly in units.
  {
    const int64_t per_cta = (int64_t)p.batch * units_row * stage_bytes * p.hsplit / grid;
    const int auto_div = per_cta < (1 << 20) ? 0 : (per_cta < (2 << 20) ? 2 : 4);  // 0: even static cut in units
    p.share_div = tuning(kTuneShareDiv) > 0 ? (tuning(kTuneShareDiv) == 99 ? 0 : tuning(kTuneShareDiv)) : auto_div;
  }
  p.no_math = tuning(kTuneNoMath) == 1;
  p.evict_first = tuning(kTuneNoEvictFirst) != 1;
  p.poll_first_ns = tuning(kTunePollFirstNs) > 0 ? tuning(kTunePollFirstNs) : 4000;
  const size_t smem = score_smem_layout(p.stages, p.hq, p.hkv, p.hsplit, p.batch, FUSED, qsplit ? 2 : 1).total;
  // no more CTAs than pieces of work: chunks, or single units when the rows are cut evenly
  const int64_t max_chunks = (int64_t)p.batch * (p.share_div == 0 ? units_row : (units_row + p.chunk_units - 1) / p.chunk_units);
  if ((int64_t)grid > max_chunks) grid = (int)max_chunks;
  if (grid < 1) grid = 1;
  if constexpr (AHPC > 0) {
    // the attention phase lives in the ring + q area of the scoring phase (everything below the barriers)
    const int region = score_smem_layout(p.stages, p.hq, p.hkv, p.hsplit, p.batch, FUSED, qsplit ? 2 : 1).bar_off;
    if (int rc = attn_configure_fused(p.attn, AHPC, region, grid)) return rc;
  }
  const cudaError_t e = qsplit ? launch_score_kernel<NT, FUSED, true, AHPC>(p, grid, smem, stream)
                               : launch_score_kernel<NT, FUSED, false, AHPC>(p, grid, smem, stream);
  if (e != cudaSuccess) return cuda_error("quest_page_score launch", e);
  return check_launch(AHPC > 0 ? "quest_decode_layer" : (FUSED ? "quest_score_select_pack" : "quest_page_score"));
}

static ScoreParams make_score_params(const void* q, int64_t q_stride_b, const void* bounds,
                                     int64_t num_chunks, const int32_t* unit_table, int64_t ut_stride,
                                     const int32_t* req_rows, const int32_t* context_lens,
                                     void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                     int n_cols) {
  ScoreParams p{};
  p.q = (const __nv_bfloat16*)q; p.q_stride_b = q_stride_b;
  p.bounds = (const __nv_bfloat16*)bounds; p.num_chunks = num_chunks;
  p.unit_table = unit_table; p.ut_stride = ut_stride; p.req_rows = req_rows;
  p.context_lens = context_lens; p.scores = (uint16_t*)scores; p.score_stride = score_stride;
  p.batch = batch; p.hq = hq; p.hkv = hkv; p.n_cols = n_cols;
  p.dbg_times = debug_times_buffer();
  return p;
}

int launch_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_chunks,
                            const int32_t* unit_table, int64_t ut_stride, const int32_t* req_rows,
                            const int32_t* context_lens, void* scores, int64_t score_stride,
                            int batch, int hq, int hkv, int n_cols, int32_t* sched, cudaStream_t stream) {
  if (batch <= 0 || n_cols <= 0) return 0;
  ScoreParams p = make_score_params(q, q_stride_b, bounds, num_chunks, unit_table, ut_stride, req_rows,
                                    context_lens, scores, score_stride, batch, hq, hkv, n_cols);
  p.sched = sched;
  return hq / hkv <= 8 ? launch_score_variant<1, false>(p, stream)
                       : launch_score_variant<2, false>(p, stream);
}

int launch_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds,
                                   int64_t num_chunks, const int32_t* unit_table, int64_t ut_stride,
                                   const int32_t* page_table, int64_t pt_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                   int n_cols, int k, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   const FusedStore* store, cudaStream_t stream) {
  if (batch <= 0) return 0;
  if (n_cols <= 0) {
    // nothing to score: the step's rows must still reach the cache
    if (store == nullptr) return 0;
    return launch_store_kv_minmax(store->key, store->k_stride, store->value, store->v_stride, store->k_cache,
                                  store->v_cache, store->slot_mapping, batch, hkv, store->num_slots,
                                  store->refresh_bounds ? const_cast<void*>(bounds) : nullptr, store->page_rec,
                                  num_chunks, stream);
  }
  ScoreParams p = make_score_params(q, q_stride_b, bounds, num_chunks, unit_table, ut_stride, req_rows,
                                    context_lens, scores, score_stride, batch, hq, hkv, n_cols);
  p.page_table = page_table; p.pt_stride = pt_stride;
  p.k = k; p.packed_slots = packed_slots; p.out_stride = out_stride; p.sparse_lens = sparse_lens;
  p.selected_pages = selected_pages; p.sched = tickets;
  if (store != nullptr) {
    p.store = StoreArgs{(const __nv_bfloat16*)store->key, store->k_stride, (const __nv_bfloat16*)store->value,
                        store->v_stride, (__nv_bfloat16*)store->k_cache, (__nv_bfloat16*)store->v_cache,
                        store->slot_mapping, store->num_slots,
                        store->refresh_bounds ? (__nv_bfloat16*)const_cast<void*>(bounds) : nullptr,
                        store->page_rec, num_chunks * kPage};
    p.store_tokens = batch;
  }
  return hq / hkv <= 8 ? launch_score_variant<1, true>(p, stream)
                       : launch_score_variant<2, true>(p, stream);
}

// One launch per sparse decode layer: (store +) score + select + pack + sparse decode attention.
int launch_quest_decode_layer(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_chunks,
                              const int32_t* unit_table, int64_t ut_stride, const int32_t* page_table,
                              int64_t pt_stride, const int32_t* req_rows, const int32_t* context_lens,
                              void* scores, int64_t score_stride, int batch, int hq, int hkv, int n_cols, int k,
                              int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens,
                              int32_t* selected_pages, int32_t* sched, const FusedStore* store,
                              const AttnParams& attn, cudaStream_t stream) {
  if (batch <= 0) return 0;
  if (n_cols <= 0 || hq / hkv > 8) {
    set_error("quest_decode_layer: needs n_cols > 0 and a group size <= 8 (got n_cols=%d, %d q heads per KV head): "
              "use the separate entry points", n_cols, hq / hkv);
    return 2;
  }
  ScoreParams p = make_score_params(q, q_stride_b, bounds, num_chunks, unit_table, ut_stride, req_rows,
                                    context_lens, scores, score_stride, batch, hq, hkv, n_cols);
  p.page_table = page_table; p.pt_stride = pt_stride;
  p.k = k; p.packed_slots = packed_slots; p.out_stride = out_stride; p.sparse_lens = sparse_lens;
  p.selected_pages = selected_pages; p.sched = sched;
  if (store != nullptr) {
    p.store = StoreArgs{(const __nv_bfloat16*)store->key, store->k_stride, (const __nv_bfloat16*)store->value,
                        store->v_stride, (__nv_bfloat16*)store->k_cache, (__nv_bfloat16*)store->v_cache,
                        store->slot_mapping, store->num_slots,
                        store->refresh_bounds ? (__nv_bfloat16*)const_cast<void*>(bounds) : nullptr,
                        store->page_rec, num_chunks * kPage};
    p.store_tokens = batch;
  }
  p.attn = attn;
  p.attn.dbg_times = nullptr;  // (the layer kernel's own time stamps carry the attention phase's: slots 16..)
  switch (attn_auto_heads_per_cta(hkv, batch, attn.max_len)) {
    case 1: return launch_score_variant<1, true, 1>(p, stream);
    case 2: return launch_score_variant<1, true, 2>(p, stream);
    default: return launch_score_variant<1, true, 4>(p, stream);
  }
}

}  // namespace svb200
