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

// bf16 bit pattern -> 16-bit key that orders like the float value.
__device__ __forceinline__ uint32_t bf16_order_key(uint32_t bits) {
  return (bits & 0x8000u) ? (~bits & 0xffffu) : (bits | 0x8000u);
}

__device__ __forceinline__ int block_exclusive_scan(int v, int* warp_scan) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads();  // warp_scan may still be read from a previous call
  if (lane == 31) warp_scan[warp] = inc;
  __syncthreads();
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
  __syncthreads();
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
__device__ void quest_select_pack_row(const uint16_t* __restrict__ scores_row, int n_cols, int k,
                                      const int32_t* __restrict__ pt_row, int ctx,
                                      int32_t* __restrict__ out_row, int32_t* __restrict__ len_out,
                                      int32_t* __restrict__ sel_pages_out, SelectSmem& sm) {
  const int tid = threadIdx.x, nthr = blockDim.x;
  const int nvec = (n_cols + 7) >> 3;
  const uint4* vrow = reinterpret_cast<const uint4*>(scores_row);

  // ---- pass 1: histogram of the high key byte
  for (int i = tid; i < 256; i += nthr) sm.hist[i] = 0;
  __syncthreads();
  for (int v = tid; v < nvec; v += nthr) {
    const uint4 x = __ldcg(vrow + v);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (v * 8 + i < n_cols) {
        const uint32_t key = bf16_order_key((w[i >> 1] >> ((i & 1) * 16)) & 0xffffu);
        atomicAdd(&sm.hist[key >> 8], 1);
      }
    }
  }
  __syncthreads();
  if (tid < 32) find_kth_bin(sm.hist, k, sm.bcast);
  __syncthreads();
  const int bin1 = sm.bcast[0];
  const int k1 = k - sm.bcast[1];
  __syncthreads();

  // ---- pass 2: histogram of the low byte inside bin1
  for (int i = tid; i < 256; i += nthr) sm.hist[i] = 0;
  __syncthreads();
  for (int v = tid; v < nvec; v += nthr) {
    const uint4 x = __ldcg(vrow + v);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (v * 8 + i < n_cols) {
        const uint32_t key = bf16_order_key((w[i >> 1] >> ((i & 1) * 16)) & 0xffffu);
        if ((int)(key >> 8) == bin1) atomicAdd(&sm.hist[key & 0xff], 1);
      }
    }
  }
  __syncthreads();
  if (tid < 32) find_kth_bin(sm.hist, k1, sm.bcast);
  __syncthreads();
  const uint32_t thresh = ((uint32_t)bin1 << 8) | (uint32_t)sm.bcast[0];
  const int need_eq = k1 - sm.bcast[1];  // how many keys == thresh to take (lowest index first)
  __syncthreads();

  // ---- pass 3: ordered compaction over contiguous per-thread chunks
  const int vec_per_thr = (nvec + nthr - 1) / nthr;
  const int v_begin = min(tid * vec_per_thr, nvec), v_end = min(v_begin + vec_per_thr, nvec);
  int cnt_gt = 0, cnt_eq = 0;
  for (int v = v_begin; v < v_end; ++v) {
    const uint4 x = __ldcg(vrow + v);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (v * 8 + i < n_cols) {
        const uint32_t key = bf16_order_key((w[i >> 1] >> ((i & 1) * 16)) & 0xffffu);
        cnt_gt += key > thresh;
        cnt_eq += key == thresh;
      }
    }
  }
  const int eq_before = block_exclusive_scan(cnt_eq, sm.warp_scan);
  const int eq_take = max(0, min(cnt_eq, need_eq - eq_before));
  int pos = block_exclusive_scan(cnt_gt + eq_take, sm.warp_scan);
  int eq_left = eq_take;
  for (int v = v_begin; v < v_end; ++v) {
    const uint4 x = __ldcg(vrow + v);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int col = v * 8 + i;
      if (col < n_cols) {
        const uint32_t key = bf16_order_key((w[i >> 1] >> ((i & 1) * 16)) & 0xffffu);
        bool take = key > thresh;
        if (key == thresh && eq_left > 0) {
          take = true;
          --eq_left;
        }
        if (take) sm.sel[pos++] = col;
      }
    }
  }
  // last page (quest.py:1888-1889: safe_num_pages - 1)
  const int num_pages = (ctx + kPage - 1) / kPage;
  if (tid == 0) sm.sel[k] = max(num_pages, 1) - 1;
  __syncthreads();
  for (int j = tid; j <= k; j += nthr) {
    const int page = sm.sel[j];
    sm.sel_slot[j] = pt_row[page];
    if (sel_pages_out) sel_pages_out[j] = page;
  }
  __syncthreads();
  // packed token slots: (k+1)*16 int32 = (k+1)*4 int4
  int4* out4 = reinterpret_cast<int4*>(out_row);
  for (int idx = tid; idx < (k + 1) * 4; idx += nthr) {
    const int base = sm.sel_slot[idx >> 2] * kPage + (idx & 3) * 4;
    out4[idx] = make_int4(base, base + 1, base + 2, base + 3);
  }
  if (tid == 0) *len_out = k * kPage + (ctx - (num_pages - 1) * kPage);
  __syncthreads();
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
  quest_select_pack_row(scores + (int64_t)b * score_stride, n_cols, k,
                        page_table + row * pt_stride, ctx, out_row, sparse_lens + b,
                        selected_pages ? selected_pages + (int64_t)b * (k + 1) : nullptr, sm);
}

// =========================================================================
// Page scoring.  One warp scores a tile of 16 logical pages for every local KV
// head; A = page bounds (16 pages x 16 k), B = q+^T / q-^T (16 k x 8 heads).
// The contraction index is permuted so each lane feeds two MMAs from ONE
// 16-byte global load of a bounds row ("k-slot permutation").
// =========================================================================
constexpr int kScoreWarps = 4;
constexpr int kScorePagesPerCta = kScoreWarps * 16;

template <bool FUSED_SELECT>
__global__ void __launch_bounds__(kScoreWarps * 32, 4) quest_page_score_kernel(
    const __nv_bfloat16* __restrict__ q, int64_t q_stride_b,
    const __nv_bfloat16* __restrict__ bounds, const int32_t* __restrict__ page_table,
    int64_t pt_stride, const int32_t* __restrict__ req_rows,
    const int32_t* __restrict__ context_lens, uint16_t* __restrict__ scores, int64_t score_stride,
    int hq, int hkv, int n_cols,
    // fused select only:
    int k, int32_t* __restrict__ packed_slots, int64_t out_stride, int32_t* __restrict__ sparse_lens,
    int32_t* __restrict__ selected_pages, int32_t* __restrict__ tickets) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  // sq[2][hq][128] bf16, 16-byte chunks XOR-swizzled by (head & 1) << 2
  uint4* sq_pos = reinterpret_cast<uint4*>(smem_raw);
  uint4* sq_neg = sq_pos + hq * 16;

  const int b = blockIdx.y;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int gsz = hq / hkv;
  const int ntiles = (gsz + 7) >> 3;

  // ---- stage q+ / q- (clamp_min(0) / clamp_max(0), quest.py:1796-1797)
  {
    const uint4* qrow = reinterpret_cast<const uint4*>(q + (int64_t)b * q_stride_b);
    for (int c = tid; c < hq * 16; c += blockDim.x) {
      const int head = c >> 4, chunk = c & 15;
      const uint4 x = qrow[c];
      uint4 p, n;
      const uint32_t w[4] = {x.x, x.y, x.z, x.w};
      uint32_t pw[4], nw[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t lo = w[i] & 0xffffu, hi = w[i] >> 16;
        const uint32_t lo_neg = (lo & 0x8000u) ? lo : 0u, lo_pos = (lo & 0x8000u) ? 0u : lo;
        const uint32_t hi_neg = (hi & 0x8000u) ? hi : 0u, hi_pos = (hi & 0x8000u) ? 0u : hi;
        pw[i] = lo_pos | (hi_pos << 16);
        nw[i] = lo_neg | (hi_neg << 16);
      }
      p = make_uint4(pw[0], pw[1], pw[2], pw[3]);
      n = make_uint4(nw[0], nw[1], nw[2], nw[3]);
      const int sw = head * 16 + (chunk ^ ((head & 1) << 2));
      sq_pos[sw] = p;
      sq_neg[sw] = n;
    }
  }
  __syncthreads();

  const int ctx = context_lens[b];
  const int n_valid = max((ctx + kPage - 1) / kPage, 1) - 1;  // valid_prev, quest.py:1883-1884
  const int64_t row = req_rows[b];
  const int p0 = blockIdx.x * kScorePagesPerCta + warp * 16 + g, p1 = p0 + 8;
  const bool v0 = p0 < n_valid && p0 < n_cols, v1 = p1 < n_valid && p1 < n_cols;
  float best0 = -INFINITY, best1 = -INFINITY;

  if (__any_sync(0xffffffffu, v0 || v1)) {
    const int32_t* pt_row = page_table + row * pt_stride;
    const int s0 = v0 ? max(pt_row[p0], 0) : 0;
    const int s1 = v1 ? max(pt_row[p1], 0) : 0;
    const int64_t page_elems = (int64_t)2 * hkv * kHeadDim;
    const __nv_bfloat16* base0 = bounds + s0 * page_elems + t * 8;
    const __nv_bfloat16* base1 = bounds + s1 * page_elems + t * 8;
    const int64_t min_off = (int64_t)hkv * kHeadDim;
    const uint4 zero4 = make_uint4(0, 0, 0, 0);

    for (int h = 0; h < hkv; ++h) {
      uint4 amax0[4], amax1[4], amin0[4], amin1[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int off = h * kHeadDim + j * 32;
        amax0[j] = v0 ? ldg_stream_16(base0 + off) : zero4;
        amin0[j] = v0 ? ldg_stream_16(base0 + min_off + off) : zero4;
        amax1[j] = v1 ? ldg_stream_16(base1 + off) : zero4;
        amin1[j] = v1 ? ldg_stream_16(base1 + min_off + off) : zero4;
      }
      for (int nt = 0; nt < ntiles; ++nt) {
        const int hg = nt * 8 + g;  // head (within the group) this lane feeds as B column
        const bool hv = hg < gsz;
        const int qh = h * gsz + (hv ? hg : 0);
        float cp[4] = {0.f, 0.f, 0.f, 0.f}, cn[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int sw = qh * 16 + ((j * 4 + t) ^ ((qh & 1) << 2));
          const uint4 bp = hv ? sq_pos[sw] : zero4;
          const uint4 bn = hv ? sq_neg[sw] : zero4;
          {
            const uint32_t a[4] = {amax0[j].x, amax1[j].x, amax0[j].y, amax1[j].y};
            mma_16816(cp, a, bp.x, bp.y);
          }
          {
            const uint32_t a[4] = {amax0[j].z, amax1[j].z, amax0[j].w, amax1[j].w};
            mma_16816(cp, a, bp.z, bp.w);
          }
          {
            const uint32_t a[4] = {amin0[j].x, amin1[j].x, amin0[j].y, amin1[j].y};
            mma_16816(cn, a, bn.x, bn.y);
          }
          {
            const uint32_t a[4] = {amin0[j].z, amin1[j].z, amin0[j].w, amin1[j].w};
            mma_16816(cn, a, bn.z, bn.w);
          }
        }
        // c0:(page g, head 2t) c1:(g, 2t+1) c2:(g+8, 2t) c3:(g+8, 2t+1)
        const int col0 = nt * 8 + 2 * t;
        float s[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) s[i] = round_bf16(round_bf16(cp[i]) + round_bf16(cn[i]));
        if (col0 < gsz) {
          best0 = fmaxf(best0, s[0]);
          best1 = fmaxf(best1, s[2]);
        }
        if (col0 + 1 < gsz) {
          best0 = fmaxf(best0, s[1]);
          best1 = fmaxf(best1, s[3]);
        }
      }
    }
    best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 1));
    best0 = fmaxf(best0, __shfl_xor_sync(0xffffffffu, best0, 2));
    best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 1));
    best1 = fmaxf(best1, __shfl_xor_sync(0xffffffffu, best1, 2));
  }
  if (t == 0) {
    uint16_t* srow = scores + (int64_t)b * score_stride;
    const uint16_t ninf = 0xff80u;
    if (p0 < n_cols) {
      __nv_bfloat16 v = __float2bfloat16_rn(best0);
      srow[p0] = v0 ? *reinterpret_cast<uint16_t*>(&v) : ninf;
    }
    if (p1 < n_cols) {
      __nv_bfloat16 v = __float2bfloat16_rn(best1);
      srow[p1] = v1 ? *reinterpret_cast<uint16_t*>(&v) : ninf;
    }
  }

  if constexpr (FUSED_SELECT) {
    __shared__ int s_last;
    __threadfence();
    __syncthreads();
    if (tid == 0) {
      const int ticket = atomicAdd(&tickets[b], 1);
      s_last = (ticket == (int)gridDim.x - 1);
      if (s_last) tickets[b] = 0;  // leave the scratch zeroed for the next launch / graph replay
    }
    __syncthreads();
    if (s_last) {
      __threadfence();
      __shared__ SelectSmem sm;
      quest_select_pack_row(scores + (int64_t)b * score_stride, n_cols, k,
                            page_table + row * pt_stride, ctx,
                            packed_slots + (int64_t)b * out_stride, sparse_lens + b,
                            selected_pages ? selected_pages + (int64_t)b * (k + 1) : nullptr, sm);
    }
  }
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

int launch_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds,
                            const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
                            const int32_t* context_lens, void* scores, int64_t score_stride,
                            int batch, int hq, int hkv, int n_cols, cudaStream_t stream) {
  if (batch <= 0 || n_cols <= 0) return 0;
  dim3 grid((n_cols + kScorePagesPerCta - 1) / kScorePagesPerCta, batch);
  const size_t smem = (size_t)2 * hq * kRowBytes;
  auto kern = quest_page_score_kernel<false>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_error("quest_page_score smem attr", e);
  }
  kern<<<grid, kScoreWarps * 32, smem, stream>>>(
      (const __nv_bfloat16*)q, q_stride_b, (const __nv_bfloat16*)bounds, page_table, pt_stride,
      req_rows, context_lens, (uint16_t*)scores, score_stride, hq, hkv, n_cols, 0, nullptr, 0,
      nullptr, nullptr, nullptr);
  return check_launch("quest_page_score");
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

int launch_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds,
                                   const int32_t* page_table, int64_t pt_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                   int n_cols, int k, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   cudaStream_t stream) {
  if (batch <= 0 || n_cols <= 0) return 0;
  dim3 grid((n_cols + kScorePagesPerCta - 1) / kScorePagesPerCta, batch);
  const size_t smem = (size_t)2 * hq * kRowBytes;
  auto kern = quest_page_score_kernel<true>;
  if (smem + sizeof(SelectSmem) > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_error("quest_score_select_pack smem attr", e);
  }
  kern<<<grid, kScoreWarps * 32, smem, stream>>>(
      (const __nv_bfloat16*)q, q_stride_b, (const __nv_bfloat16*)bounds, page_table, pt_stride,
      req_rows, context_lens, (uint16_t*)scores, score_stride, hq, hkv, n_cols, k, packed_slots,
      out_stride, sparse_lens, selected_pages, tickets);
  return check_launch("quest_score_select_pack");
}

}  // namespace svb200
