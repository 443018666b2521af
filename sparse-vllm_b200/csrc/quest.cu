// Quest kernels for sm_100a: page min/max bounds, fused gather-free page scoring,
// radix top-k page selection + slot packing.
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/engine/cache_manager/quest.py):
//   bounds  : on_kv_stored 1608-1716, on_forward_end 1719-1771
//   scoring : _score_pages_batched 1773-1802 (three bf16 roundings per head score)
//   select  : _build_decode_view_static 1833-1913
#include "common.cuh"
#include "launch.h"

namespace svb200 {

// =========================================================================
// Page bounds.  bounds[page][2][Hkv][128]; plane 0 = max, plane 1 = min.
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
    const int32_t* __restrict__ slot_mapping, int n_tokens, int hkv, int64_t num_slots) {
  const int lane = threadIdx.x & 31;
  const int warps_per_block = blockDim.x >> 5;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const int vecs = hkv * (kHeadDim / 8);
  for (int tok = blockIdx.x * warps_per_block + (threadIdx.x >> 5); tok < n_tokens;
       tok += gridDim.x * warps_per_block) {
    const int slot = slot_mapping[tok];
    if (slot < 0 || (slot & (kPage - 1)) != kPage - 1 || slot >= num_slots) continue;
    const int64_t page = slot >> 4;
    const __nv_bfloat16* page_base = k_cache + page * kPage * row_elems;
    __nv_bfloat16* out = bounds + page * 2 * row_elems;
    for (int v = lane; v < vecs; v += 32) {
      uint4 mx, mn;
      page_minmax_vec(page_base + v * 8, row_elems, mx, mn);
      *reinterpret_cast<uint4*>(out + v * 8) = mx;
      *reinterpret_cast<uint4*>(out + row_elems + v * 8) = mn;
    }
  }
}

// grid (batch, layers): rebuild the bounds of the page a decode token just completed.
__global__ void __launch_bounds__(128) quest_decode_page_complete_kernel(
    const __nv_bfloat16* __restrict__ k_cache_all, __nv_bfloat16* __restrict__ bounds_all,
    const int32_t* __restrict__ page_table, int64_t pt_stride,
    const int32_t* __restrict__ req_rows, const int32_t* __restrict__ seq_lens, int hkv,
    int64_t num_slots, int64_t num_pages) {
  const int b = blockIdx.x, layer = blockIdx.y;
  const int len = seq_lens[b];
  if (len <= 0 || (len & (kPage - 1)) != 0) return;
  const int64_t row = req_rows[b];
  const int slot_page = page_table[row * pt_stride + (len / kPage - 1)];
  if (slot_page < 0 || slot_page >= num_pages) return;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const __nv_bfloat16* page_base =
      k_cache_all + ((int64_t)layer * num_slots + (int64_t)slot_page * kPage) * row_elems;
  __nv_bfloat16* out = bounds_all + ((int64_t)layer * num_pages + slot_page) * 2 * row_elems;
  const int vecs = hkv * (kHeadDim / 8);
  for (int v = threadIdx.x; v < vecs; v += blockDim.x) {
    uint4 mx, mn;
    page_minmax_vec(page_base + v * 8, row_elems, mx, mn);
    *reinterpret_cast<uint4*>(out + v * 8) = mx;
    *reinterpret_cast<uint4*>(out + row_elems + v * 8) = mn;
  }
}

// =========================================================================
// Top-k selection + packing of one row (device routine shared by the standalone
// and the fused kernel).  Works for any block size that is a multiple of 32.
// =========================================================================
constexpr int kMaxPrevBudget = 2047;

struct SelectSmem {
  int hist[256];
  int warp_scan[32];
  int bcast[4];
  int sel[kMaxPrevBudget + 1];
  int sel_slot[kMaxPrevBudget + 1];
};

constexpr uint32_t kNegInfKey = 0x007fu;  // order key of bf16 -inf (0xff80)
// bf16 bit pattern -> 16-bit key that orders like the float value.
__device__ __forceinline__ uint32_t bf16_order_key(uint32_t bits) {
  return (bits & 0x8000u) ? (~bits & 0xffffu) : (bits | 0x8000u);
}

// Barrier over the first NTHR threads of the CTA (barrier 0 = __syncthreads when NTHR is the
// whole block; the fused scoring kernel keeps its producer warp out with barrier 1).
template <int BAR, int NTHR>
__device__ __forceinline__ void group_sync() {
  asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(NTHR) : "memory");
}

template <int BAR, int NTHR>
__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_scan) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = NTHR >> 5;
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
  }
  group_sync<BAR, NTHR>();
  return warp_scan[warp] + inc - v;
}

// Finds the bin holding the k-th largest item: largest bin B with sum(hist[B..255]) >= k.
// Returns B and the count strictly above it.  Executed by warp 0; result via bcast.
__device__ __forceinline__ void find_kth_bin(const int* hist, int k, int* bcast) {
  const int lane = threadIdx.x & 31;
  // lane L owns bins [8L, 8L+8); higher lanes = higher bins.
  int local[8];
  int lsum = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    local[i] = hist[lane * 8 + i];
    lsum += local[i];
  }
  // suffix sum over lanes (inclusive)
  int suf = lsum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    int n = __shfl_down_sync(0xffffffffu, suf, o);
    if (lane + o < 32) suf += n;
  }
  const int above_lane = suf - lsum;  // items in bins of higher lanes
  // the owning lane: above_lane < k <= suf
  if (above_lane < k && k <= suf) {
    int above = above_lane;
    int bin = lane * 8 + 7;
#pragma unroll
    for (int i = 7; i >= 0; --i) {
      if (above + local[i] >= k) {
        bin = lane * 8 + i;
        break;
      }
      above += local[i];
    }
    bcast[0] = bin;
    bcast[1] = above;
  }
}

// scores_row: bf16 bits, 16-byte aligned, readable up to round_up(n_cols, 8).
// Columns >= n_valid count as -inf whatever the buffer holds (quest.py:1883-1885).
// Three passes over the row (high-byte histogram, low-byte histogram, ordered compaction); each
// thread owns a contiguous chunk of 8-score vectors.  CACHED keeps the chunk in registers so the
// row is read from L2 once (rows up to 4 * NTHR vectors = 8192 scores at 256 threads).
constexpr int kSelectVecPerThread = 4;

template <int BAR, int NTHR, bool CACHED>
__device__ void quest_select_pack_impl(const uint16_t* __restrict__ scores_row, int n_cols, int n_valid,
                                       int k, const int32_t* __restrict__ pt_row, int ctx,
                                       int32_t* __restrict__ out_row, int32_t* __restrict__ len_out,
                                       int32_t* __restrict__ sel_pages_out, SelectSmem& sm) {
  constexpr int VPT = kSelectVecPerThread;
  const int tid = threadIdx.x;
  const int nvec = (n_cols + 7) >> 3;
  const uint4* vrow = reinterpret_cast<const uint4*>(scores_row);
  const int vec_per_thr = CACHED ? VPT : (nvec + NTHR - 1) / NTHR;
  const int v_begin = min(tid * vec_per_thr, nvec), v_end = min(v_begin + vec_per_thr, nvec);
  uint4 xr[VPT];
  if constexpr (CACHED) {
#pragma unroll
    for (int vv = 0; vv < VPT; ++vv)
      xr[vv] = v_begin + vv < v_end ? __ldcg(vrow + v_begin + vv) : make_uint4(0, 0, 0, 0);
  }
  // body(col, key, valid) for every column slot of this thread's chunk, in ascending column order.
  // Control flow is warp-uniform (invalid slots are passed with valid = false) so bodies may use
  // full-mask warp primitives.
  auto for_each_key = [&](auto&& body) {
    auto one_vec = [&](int v, bool vec_ok, const uint4& x) {
      const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int col = v * 8 + i;
        body(col, col < n_valid ? bf16_order_key((w[i >> 1] >> ((i & 1) * 16)) & 0xffffu) : kNegInfKey,
             vec_ok && col < n_cols);
      }
    };
    if constexpr (CACHED) {
#pragma unroll
      for (int vv = 0; vv < VPT; ++vv) one_vec(v_begin + vv, v_begin + vv < v_end, xr[vv]);
    } else {
      for (int vv = 0; vv < vec_per_thr; ++vv) {
        const bool ok = v_begin + vv < v_end;
        one_vec(v_begin + vv, ok, ok ? __ldcg(vrow + v_begin + vv) : make_uint4(0, 0, 0, 0));
      }
    }
  };
  // Histogram increment aggregated over the lanes that hit the same bin: page scores crowd into a
  // few exponent bins, and one shared-memory atomic per lane would serialise on them.
  auto hist_add = [&](int bin, bool valid) {
    const unsigned peers = __match_any_sync(0xffffffffu, valid ? bin : -1);
    if (valid && (tid & 31) == __ffs(peers) - 1) atomicAdd(&sm.hist[bin], __popc(peers));
  };

  // ---- pass 1: histogram of the high key byte
  for (int i = tid; i < 256; i += NTHR) sm.hist[i] = 0;
  group_sync<BAR, NTHR>();
  for_each_key([&](int, uint32_t key, bool valid) { hist_add(key >> 8, valid); });
  group_sync<BAR, NTHR>();
  if (tid < 32) find_kth_bin(sm.hist, k, sm.bcast);
  group_sync<BAR, NTHR>();
  const int bin1 = sm.bcast[0];
  const int k1 = k - sm.bcast[1];
  group_sync<BAR, NTHR>();

  // ---- pass 2: histogram of the low byte inside bin1
  for (int i = tid; i < 256; i += NTHR) sm.hist[i] = 0;
  group_sync<BAR, NTHR>();
  for_each_key([&](int, uint32_t key, bool valid) {
    hist_add(key & 0xff, valid && (int)(key >> 8) == bin1);
  });
  group_sync<BAR, NTHR>();
  if (tid < 32) find_kth_bin(sm.hist, k1, sm.bcast);
  group_sync<BAR, NTHR>();
  const uint32_t thresh = ((uint32_t)bin1 << 8) | (uint32_t)sm.bcast[0];
  const int need_eq = k1 - sm.bcast[1];  // how many keys == thresh to take (lowest index first)
  group_sync<BAR, NTHR>();

  // ---- pass 3: ordered compaction
  int cnt_gt = 0, cnt_eq = 0;
  for_each_key([&](int, uint32_t key, bool valid) {
    cnt_gt += valid && key > thresh;
    cnt_eq += valid && key == thresh;
  });
  const int eq_before = block_exclusive_scan<BAR, NTHR>(cnt_eq, sm.warp_scan);
  const int eq_take = max(0, min(cnt_eq, need_eq - eq_before));
  int pos = block_exclusive_scan<BAR, NTHR>(cnt_gt + eq_take, sm.warp_scan);
  int eq_left = eq_take;
  for_each_key([&](int col, uint32_t key, bool valid) {
    bool take = valid && key > thresh;
    if (valid && key == thresh && eq_left > 0) {
      take = true;
      --eq_left;
    }
    if (take) sm.sel[pos++] = col;
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
}

template <int BAR, int NTHR>
__device__ void quest_select_pack_row(const uint16_t* __restrict__ scores_row, int n_cols, int n_valid, int k,
                                      const int32_t* __restrict__ pt_row, int ctx,
                                      int32_t* __restrict__ out_row, int32_t* __restrict__ len_out,
                                      int32_t* __restrict__ sel_pages_out, SelectSmem& sm) {
  if (((n_cols + 7) >> 3) <= kSelectVecPerThread * NTHR)
    quest_select_pack_impl<BAR, NTHR, true>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out,
                                            sel_pages_out, sm);
  else
    quest_select_pack_impl<BAR, NTHR, false>(scores_row, n_cols, n_valid, k, pt_row, ctx, out_row, len_out,
                                             sel_pages_out, sm);
}

__global__ void __launch_bounds__(256) quest_select_pack_kernel(
    const uint16_t* __restrict__ scores, int64_t score_stride, int n_cols,
    const int32_t* __restrict__ page_table, int64_t pt_stride, const int32_t* __restrict__ req_rows,
    const int32_t* __restrict__ context_lens, int k, int32_t* __restrict__ packed_slots,
    int64_t out_stride, int32_t* __restrict__ sparse_lens, int32_t* __restrict__ selected_pages,
    const int32_t* __restrict__ dense_table, int64_t dense_stride, int token_budget,
    int page_budget_base, int out_width) {
  __shared__ SelectSmem sm;
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
// Stream-K over page tiles: a "unit" is 16 consecutive logical pages of one batch
// row; only valid pages (p < ceil(ctx/16) - 1) form units.  The launch's units are
// cut into one equal contiguous range per persistent CTA.  Warps 8-11 (producers)
// resolve 16 page slots through the page table and copy each page's bounds record
// ([2][Hkv][128] = Hkv*512 contiguous bytes) with 16-byte LDGSTS, a warp per 512 B,
// into a ring of S stages (130-200 KB in flight per SM) whose mbarriers complete
// when the copies land.  Warps 0-7
// (consumers) take units round-robin; per unit and KV head:
//   A = bounds rows (16 pages x 128, ldmatrix), B = q+ / q- (128 x 8 heads)
//   score = bf16(bf16(A_max.q+) + bf16(A_min.q-)), then max over all local heads.
// When the last contributor of a row finishes (arrival ticket), its consumer
// warps run the top-k selection and slot packing for that row in the same launch.
// =========================================================================
constexpr int kScoreConsumerWarps = 8;
constexpr int kScoreConsumerThreads = kScoreConsumerWarps * 32;
constexpr int kScoreProducerWarps = 4;
constexpr int kPagePrefetch = 6;
constexpr int kScoreThreads = kScoreConsumerThreads + kScoreProducerWarps * 32;
constexpr int kScoreMaxSmem = 227 * 1024;

struct ScoreParams {
  const __nv_bfloat16* q;
  int64_t q_stride_b;
  const __nv_bfloat16* bounds;
  int64_t num_pages;
  const int32_t* page_table;
  int64_t pt_stride;
  const int32_t* req_rows;
  const int32_t* context_lens;
  uint16_t* scores;
  int64_t score_stride;
  int batch, hq, hkv, n_cols, stages;
  unsigned long long* dbg_times;  // experiments: per-CTA timestamps (svb200_debug_times), else null
  int consumers;  // consumer warps that take units; divides `stages` (see the kernel)
  int producers;  // producer warps that feed units; divides `stages`
  // fused select only
  int k;
  int32_t* packed_slots;
  int64_t out_stride;
  int32_t* sparse_lens;
  int32_t* selected_pages;
  int32_t* tickets;
};

struct ScoreSmemLayout {
  int ring_off, q_off, bar_off, valid_off, prefix_off, select_off, total;
};

__host__ __device__ inline int score_rec_stride(int hkv) { return hkv * 2 * kRowBytes + 16; }

__host__ __device__ inline ScoreSmemLayout score_smem_layout(int stages, int hq, int hkv, int batch,
                                                             bool fused) {
  ScoreSmemLayout L;
  int off = 0;
  L.ring_off = off;   off += stages * kPage * score_rec_stride(hkv);
  L.q_off = off;      off += 2 * hq * kRowBytes;
  L.bar_off = off;    off += 2 * stages * 8;
  L.valid_off = off;  off += ((stages * 4 + 15) / 16) * 16;
  L.prefix_off = off; off += ((batch + 1) * 4 + 15) / 16 * 16;
  L.select_off = off; off += fused ? (int)((sizeof(SelectSmem) + 15) / 16 * 16) : 0;
  L.total = off;
  return L;
}

__device__ __forceinline__ void score_consumer_sync() { group_sync<1, kScoreConsumerThreads>(); }

__device__ __forceinline__ int score_unit_owner(int64_t u, int64_t U, int64_t G) {
  return (int)(((u + 1) * G + U - 1) / U - 1);
}

template <int NT, bool FUSED_SELECT>
__global__ void __launch_bounds__(kScoreThreads, 1) quest_page_score_kernel(ScoreParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int S = p.stages, hq = p.hq, hkv = p.hkv, B = p.batch;
  const int gsz = hq / hkv;
  const int rec_stride = score_rec_stride(hkv);
  const int rec_bytes = hkv * 2 * kRowBytes;
  const int stage_bytes = kPage * rec_stride;
  const ScoreSmemLayout L = score_smem_layout(S, hq, hkv, B, FUSED_SELECT);
  unsigned char* ring = smem + L.ring_off;
  // sq[sign][head][kk][t] : uint2 {b0, b1} B fragments, kk XOR-swizzled by head
  uint32_t* sq = reinterpret_cast<uint32_t*>(smem + L.q_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* empty_bar = full_bar + S;
  uint32_t* sValid = reinterpret_cast<uint32_t*>(smem + L.valid_off);
  int* sPrefix = reinterpret_cast<int*>(smem + L.prefix_off);
  __shared__ int s_flag;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long* dbg = p.dbg_times ? p.dbg_times + (size_t)blockIdx.x * 8 : nullptr;
  if (dbg && tid == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    dbg[0] = global_timer_ns();
    dbg[7] = smid;
    dbg[5] = 0;
  }

  // ---- setup
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 33);  // 32 async copy arrivals + the producer's release
      mbar_init(&empty_bar[s], 1);
    }
    mbar_fence_init();
  }
  if (warp == 0) {  // units per row -> exclusive prefix
    int carry = 0;
    for (int base = 0; base < B; base += 32) {
      const int b = base + lane;
      int n = 0;
      if (b < B) {
        const int ctx = p.context_lens[b];
        const int nv = min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);
        n = (nv + 15) >> 4;
      }
      int inc = n;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
      }
      if (b < B) sPrefix[b] = carry + inc - n;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) sPrefix[B] = carry;
  }
  // columns no unit covers are -inf (quest.py:1883-1885)
  for (int b = blockIdx.x; b < B; b += gridDim.x) {
    const int ctx = p.context_lens[b];
    const int nv = min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);
    uint16_t* srow = p.scores + (int64_t)b * p.score_stride;
    for (int c = ((nv + 15) & ~15) + tid; c < p.n_cols; c += kScoreThreads) srow[c] = 0xff80u;
  }
  __syncthreads();

  if (dbg && tid == 0) dbg[1] = global_timer_ns();
  const int64_t U = sPrefix[B];
  const int64_t Gn = U < (int64_t)gridDim.x ? U : (int64_t)gridDim.x;
  const int u0 = Gn > 0 && (int64_t)blockIdx.x < Gn ? (int)(U * blockIdx.x / Gn) : 0;
  const int u1 = Gn > 0 && (int64_t)blockIdx.x < Gn ? (int)(U * (blockIdx.x + 1) / Gn) : 0;
  int r_first = 0;
  if (u1 > u0) {
    int lo = 0, hi = B;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (sPrefix[mid] <= u0) lo = mid; else hi = mid;
    }
    r_first = lo;
  }

  if (warp >= kScoreConsumerWarps) {
    // ============================================================== producers
    // unit i is fed by producer i % NP, NP | S (a stage is always refilled by the same warp)
    const int n_units = u1 - u0;
    const int pw = warp - kScoreConsumerWarps;
    const int NP = p.producers;
    if (pw >= NP) return;
    const int pg_lane = lane & 15;
    int r_pf = r_first;
    auto row_nv = [&](int r) {
      const int ctx = p.context_lens[r];
      return min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);
    };
    int nv_pf = row_nv(r_pf);
    const int32_t* row_pf = p.page_table + (int64_t)p.req_rows[r_pf] * p.pt_stride;
    auto fetch = [&](int i) -> int {  // raw physical page of logical page pg_lane in local unit i
      if (i >= n_units) return -1;
      const int u = u0 + i;
      if (sPrefix[r_pf + 1] <= u) {
        while (sPrefix[r_pf + 1] <= u) ++r_pf;
        nv_pf = row_nv(r_pf);
        row_pf = p.page_table + (int64_t)p.req_rows[r_pf] * p.pt_stride;
      }
      const int col = (u - sPrefix[r_pf]) * 16 + pg_lane;
      if (col >= nv_pf) return -1;
      return __ldg(row_pf + col);
    };
    int s = pw % S;
    uint32_t ph = (pw / S) & 1;  // phase parity of stage s at unit i
    const char* src0 = reinterpret_cast<const char*>(p.bounds) + lane * 16;
    // page ids are fetched kPagePrefetch units of this warp ahead (the table comes from DRAM)
    int pf[kPagePrefetch];
#pragma unroll
    for (int d = 0; d < kPagePrefetch; ++d) pf[d] = fetch(pw + d * NP);
    for (int i = pw; i < n_units; i += NP) {
      const int page_new = fetch(i + kPagePrefetch * NP);
      const int pa = pf[0];
      const int my = (pa >= 0 && pa < p.num_pages) ? pa : -1;
      const uint32_t mask = __ballot_sync(0xffffffffu, my >= 0) & 0xffffu;
      const uint32_t dst0 = smem_u32(ring + s * stage_bytes) + lane * 16;
      mbar_wait(&empty_bar[s], ph ^ 1);
#pragma unroll 4
      for (int pg = 0; pg < kPage; ++pg) {
        const int page = __shfl_sync(0xffffffffu, my, pg);
        const int nbytes = page >= 0 ? 16 : 0;
        const char* src = src0 + (int64_t)max(page, 0) * rec_bytes;
        const uint32_t dst = dst0 + pg * rec_stride;
        for (int c = 0; c < hkv; ++c) cp_async_16(dst + c * 512, src + c * 512, nbytes);
      }
      if (lane == 0) sValid[s] = mask;
      cp_async_mbar_arrive(&full_bar[s]);
      if (lane == 0) mbar_arrive(&full_bar[s]);  // release: publishes sValid
#pragma unroll
      for (int d = 0; d + 1 < kPagePrefetch; ++d) pf[d] = pf[d + 1];
      pf[kPagePrefetch - 1] = page_new;
      s += NP;
      while (s >= S) s -= S, ph ^= 1;
    }
    if (dbg && pw == 0 && lane == 0) dbg[4] = global_timer_ns();
    return;
  }

  // ================================================================ consumers
  const int g = lane >> 2, t = lane & 3;
  const uint2* sq2 = reinterpret_cast<const uint2*>(sq);
  const int neg_off = hq * 32;  // uint2 offset of the q- fragments
  [[maybe_unused]] SelectSmem* sel_sm = reinterpret_cast<SelectSmem*>(smem + L.select_off);

  if constexpr (FUSED_SELECT) {
    // rows without a single valid previous page never become a unit: select them here
    for (int b = blockIdx.x; b < B; b += gridDim.x) {
      if (sPrefix[b + 1] != sPrefix[b]) continue;
      const int ctx = p.context_lens[b];
      quest_select_pack_row<1, kScoreConsumerThreads>(
          p.scores + (int64_t)b * p.score_stride, p.n_cols, 0, p.k,
          p.page_table + (int64_t)p.req_rows[b] * p.pt_stride, ctx,
          p.packed_slots + (int64_t)b * p.out_stride, p.sparse_lens + b,
          p.selected_pages ? p.selected_pages + (int64_t)b * (p.k + 1) : nullptr, *sel_sm);
    }
  }

  int r = r_first, i = 0;
  while (u0 + i < u1) {
    while (sPrefix[r + 1] <= u0 + i) ++r;
    const int b = r;
    const int row_begin = sPrefix[r], row_end = sPrefix[r + 1];
    const int seg_begin = u0 + i, seg_end = min(u1, row_end);
    const int n = seg_end - seg_begin;
    const int ctx = p.context_lens[b];
    const int nv = min(max((ctx + kPage - 1) / kPage, 1) - 1, p.n_cols);

    // ---- q+ / q- (clamp_min(0) / clamp_max(0), quest.py:1796-1797) in B-fragment order
    {
      const uint4* qrow = reinterpret_cast<const uint4*>(p.q + (int64_t)b * p.q_stride_b);
      for (int c = tid; c < hq * 16; c += kScoreConsumerThreads) {
        const int head = c >> 4, chunk = c & 15;
        const uint4 x = qrow[c];
        const uint32_t w[4] = {x.x, x.y, x.z, x.w};
        const int kk = chunk >> 1, hf = chunk & 1;
        const int base = ((head * 8 + (kk ^ (head & 7))) * 4) * 2 + hf;
#pragma unroll
        for (int tt = 0; tt < 4; ++tt) {
          const uint32_t lo = w[tt] & 0xffffu, hi = w[tt] >> 16;
          const uint32_t lo_neg = (lo & 0x8000u) ? lo : 0u, lo_pos = (lo & 0x8000u) ? 0u : lo;
          const uint32_t hi_neg = (hi & 0x8000u) ? hi : 0u, hi_pos = (hi & 0x8000u) ? 0u : hi;
          sq[base + tt * 2] = lo_pos | (hi_pos << 16);
          sq[neg_off * 2 + base + tt * 2] = lo_neg | (hi_neg << 16);
        }
      }
    }
    score_consumer_sync();

    uint16_t* srow = p.scores + (int64_t)b * p.score_stride;
    // Unit j goes to warp j % NW, and NW divides S: a stage is always drained by the same warp, so
    // the phase a warp waits for is never more than one ahead of the barrier (parity waits cannot
    // tell a phase from the one two before it).
    const int NW = p.consumers;
    for (int j = i + ((warp - i % NW + NW) % NW); warp < NW && j < i + n; j += NW) {
      const int s = j % S;
      mbar_wait(&full_bar[s], (j / S) & 1);
      if (dbg && tid == 0 && j == 0) dbg[2] = global_timer_ns();
      const uint32_t mask = sValid[s];
      const uint32_t abase = smem_u32(ring + s * stage_bytes) +
                             ((lane & 7) + ((lane >> 3) & 1) * 8) * rec_stride + (lane >> 4) * 16;
      float best0 = -INFINITY, best1 = -INFINITY;
      for (int h = 0; h < hkv; ++h) {
        float cp[NT][4], cn[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt)
#pragma unroll
          for (int e = 0; e < 4; ++e) cp[nt][e] = cn[nt][e] = 0.f;
        const uint32_t amax_addr = abase + h * kRowBytes;
        const uint32_t amin_addr = amax_addr + hkv * kRowBytes;
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          uint32_t amax[4], amin[4];
          ldmatrix_x4(amax, amax_addr + kk * 32);
          ldmatrix_x4(amin, amin_addr + kk * 32);
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const int qh = h * gsz + min(nt * 8 + g, gsz - 1);
            const int fi = (qh * 8 + (kk ^ (qh & 7))) * 4 + t;
            const uint2 bp = sq2[fi], bn = sq2[neg_off + fi];
            mma_16816(cp[nt], amax, bp.x, bp.y);
            mma_16816(cn[nt], amin, bn.x, bn.y);
          }
        }
        // c0:(page g, head 2t) c1:(g, 2t+1) c2:(g+8, 2t) c3:(g+8, 2t+1)
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float sc[4];
#pragma unroll
          for (int e = 0; e < 4; ++e) sc[e] = round_bf16(round_bf16(cp[nt][e]) + round_bf16(cn[nt][e]));
          const int col0 = nt * 8 + 2 * t;
          if (col0 < gsz) best0 = fmaxf(best0, sc[0]), best1 = fmaxf(best1, sc[2]);
          if (col0 + 1 < gsz) best0 = fmaxf(best0, sc[1]), best1 = fmaxf(best1, sc[3]);
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
      best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 1));
      best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 2));
      best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 1));
      best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 2));
      if (t == 0) {
        const int p0 = (seg_begin + (j - i) - row_begin) * 16 + g, p1 = p0 + 8;
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

    if constexpr (FUSED_SELECT) {
      const bool complete = seg_begin == row_begin && seg_end == row_end;
      __threadfence();
      score_consumer_sync();
      if (tid == 0) {
        int last = 1;
        if (!complete) {
          const int c_first = score_unit_owner(row_begin, U, Gn);
          const int c_last = score_unit_owner(row_end - 1, U, Gn);
          const int ticket = atomicAdd(&p.tickets[b], 1);
          last = (ticket == c_last - c_first);
          if (last) p.tickets[b] = 0;  // scratch stays zeroed for the next launch / replay
        }
        s_flag = last;
      }
      score_consumer_sync();
      if (s_flag) {
        __threadfence();
        const unsigned long long ts = dbg ? global_timer_ns() : 0;
        quest_select_pack_row<1, kScoreConsumerThreads>(
            srow, p.n_cols, nv, p.k, p.page_table + (int64_t)p.req_rows[b] * p.pt_stride, ctx,
            p.packed_slots + (int64_t)b * p.out_stride, p.sparse_lens + b,
            p.selected_pages ? p.selected_pages + (int64_t)b * (p.k + 1) : nullptr, *sel_sm);
        if (dbg && tid == 0) dbg[5] += global_timer_ns() - ts;
      }
    }
    score_consumer_sync();  // sq / s_flag are reused by the next segment
    i += n;
  }
  if (dbg && tid == 0) dbg[3] = global_timer_ns();
}

// =========================================================================
// Host launchers
// =========================================================================
int launch_quest_page_minmax(const void* k_cache, void* bounds_layer, const int32_t* slot_mapping,
                             int n_tokens, int hkv, int64_t num_slots, cudaStream_t stream) {
  if (n_tokens <= 0) return 0;
  const int warps_per_block = 8;
  int blocks = (n_tokens + warps_per_block - 1) / warps_per_block;
  if (blocks > 148 * 8) blocks = 148 * 8;
  quest_page_minmax_kernel<<<blocks, warps_per_block * 32, 0, stream>>>(
      (const __nv_bfloat16*)k_cache, (__nv_bfloat16*)bounds_layer, slot_mapping, n_tokens, hkv,
      num_slots);
  return check_launch("quest_page_minmax");
}

int launch_quest_decode_page_complete(const void* k_cache_all, void* bounds_all,
                                      const int32_t* page_table, int64_t pt_stride,
                                      const int32_t* req_rows, const int32_t* seq_lens, int batch,
                                      int layers, int hkv, int64_t num_slots, int64_t num_pages,
                                      cudaStream_t stream) {
  if (batch <= 0 || layers <= 0) return 0;
  dim3 grid(batch, layers);
  quest_decode_page_complete_kernel<<<grid, 128, 0, stream>>>(
      (const __nv_bfloat16*)k_cache_all, (__nv_bfloat16*)bounds_all, page_table, pt_stride,
      req_rows, seq_lens, hkv, num_slots, num_pages);
  return check_launch("quest_decode_page_complete");
}

int launch_quest_select_pack(const void* scores, int64_t score_stride, int n_cols,
                             const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
                             const int32_t* context_lens, int batch, int k, int32_t* packed_slots,
                             int64_t out_stride, int32_t* sparse_lens, int32_t* selected_pages,
                             const int32_t* dense_table, int64_t dense_stride, int token_budget,
                             int page_budget_base, int out_width, cudaStream_t stream) {
  if (batch <= 0) return 0;
  quest_select_pack_kernel<<<batch, 256, 0, stream>>>(
      (const uint16_t*)scores, score_stride, n_cols, page_table, pt_stride, req_rows, context_lens,
      k, packed_slots, out_stride, sparse_lens, selected_pages, dense_table, dense_stride,
      token_budget, page_budget_base, out_width);
  return check_launch("quest_select_pack");
}

template <int NT, bool FUSED>
static int launch_score_variant(ScoreParams p, cudaStream_t stream) {
  const int fixed = score_smem_layout(0, p.hq, p.hkv, p.batch, FUSED).total + 256;  // + static smem
  const int stage_bytes = kPage * score_rec_stride(p.hkv);
  int stages = (kScoreMaxSmem - fixed) / (stage_bytes + 20);
  if (stages > 24) stages = 24;
  if (stages < 2) {
    set_error("quest_page_score: %d q heads / %d kv heads / batch %d leave no shared memory for the bounds ring",
              p.hq, p.hkv, p.batch);
    return 1;
  }
  // (stages, consumer warps, producer warps) with consumers | stages and producers | stages:
  // prefer more producers + consumers, then more stages
  int best_s = stages, best_c = 1, best_p = 1;
  for (int s2 = stages; s2 >= 2 && s2 >= stages - 3; --s2) {
    int c = 1, pr = 1;
    for (int d = kScoreConsumerWarps; d >= 1; --d)
      if (s2 % d == 0) { c = d; break; }
    for (int d = kScoreProducerWarps; d >= 1; --d)
      if (s2 % d == 0) { pr = d; break; }
    if (c + 2 * pr > best_c + 2 * best_p || s2 == stages) best_c = c, best_p = pr, best_s = s2;
  }
  stages = best_s;
  p.stages = stages;
  p.consumers = best_c;
  p.producers = best_p;
  const size_t smem = score_smem_layout(stages, p.hq, p.hkv, p.batch, FUSED).total;
  auto kern = quest_page_score_kernel<NT, FUSED>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_error("quest_page_score smem attr", e);
  const int64_t max_units = (int64_t)p.batch * ((p.n_cols + 15) / 16);
  int grid = max_ctas();
  if ((int64_t)grid > max_units) grid = (int)max_units;
  if (FUSED && grid > p.batch && max_units < p.batch) grid = p.batch;
  if (grid < 1) grid = 1;
  kern<<<grid, kScoreThreads, smem, stream>>>(p);
  return check_launch(FUSED ? "quest_score_select_pack" : "quest_page_score");
}

static ScoreParams make_score_params(const void* q, int64_t q_stride_b, const void* bounds,
                                     int64_t num_pages, const int32_t* page_table, int64_t pt_stride,
                                     const int32_t* req_rows, const int32_t* context_lens,
                                     void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                     int n_cols) {
  ScoreParams p{};
  p.q = (const __nv_bfloat16*)q; p.q_stride_b = q_stride_b;
  p.bounds = (const __nv_bfloat16*)bounds; p.num_pages = num_pages;
  p.page_table = page_table; p.pt_stride = pt_stride; p.req_rows = req_rows;
  p.context_lens = context_lens; p.scores = (uint16_t*)scores; p.score_stride = score_stride;
  p.batch = batch; p.hq = hq; p.hkv = hkv; p.n_cols = n_cols;
  p.dbg_times = debug_times_buffer();
  return p;
}

int launch_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_pages,
                            const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
                            const int32_t* context_lens, void* scores, int64_t score_stride,
                            int batch, int hq, int hkv, int n_cols, cudaStream_t stream) {
  if (batch <= 0 || n_cols <= 0) return 0;
  ScoreParams p = make_score_params(q, q_stride_b, bounds, num_pages, page_table, pt_stride, req_rows,
                                    context_lens, scores, score_stride, batch, hq, hkv, n_cols);
  return hq / hkv <= 8 ? launch_score_variant<1, false>(p, stream)
                       : launch_score_variant<2, false>(p, stream);
}

int launch_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds,
                                   int64_t num_pages, const int32_t* page_table, int64_t pt_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                   int n_cols, int k, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   cudaStream_t stream) {
  if (batch <= 0 || n_cols <= 0) return 0;
  ScoreParams p = make_score_params(q, q_stride_b, bounds, num_pages, page_table, pt_stride, req_rows,
                                    context_lens, scores, score_stride, batch, hq, hkv, n_cols);
  p.k = k; p.packed_slots = packed_slots; p.out_stride = out_stride; p.sparse_lens = sparse_lens;
  p.selected_pages = selected_pages; p.tickets = tickets;
  return hq / hkv <= 8 ? launch_score_variant<1, true>(p, stream)
                       : launch_score_variant<2, true>(p, stream);
}

}  // namespace svb200
