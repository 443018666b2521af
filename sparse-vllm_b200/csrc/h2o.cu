// H2O kernels for sm_100a: layer-end score normalisation with fused cumulative
// accumulation, and burst eviction (heavy-hitter selection + index-only compaction).
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/engine/):
//   sparse_controller.py:761-767         score *= D^-0.5 ; softmax over the buffer width
//   cache_manager/h2o.py:958-1038        cumulative[:kv_len] = step + pad(previous, 1)
//   cache_manager/h2o.py:519-563         heavy hitters (stable descending) + recent suffix
//   cache_manager/snapkv.py:1681-1803    slot-row compaction + free-stack push
#include "common.cuh"
#include "launch.h"

namespace svb200 {

constexpr int kH2oThreads = 256;

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = lane < nwarps ? red[lane] : -INFINITY;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = lane < nwarps ? red[lane] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  return r;
}

// One CTA per row.  The scaled row is cached in shared memory when it fits.
template <bool CACHE_ROW>
__global__ void __launch_bounds__(kH2oThreads) h2o_normalize_accumulate_kernel(
    float* __restrict__ score, int64_t score_stride, int width, float scale,
    float* __restrict__ cumulative, int64_t cum_stride, const int32_t* __restrict__ cum_rows,
    const int32_t* __restrict__ kv_lens, float refill) {
  extern __shared__ float row_cache[];
  __shared__ float red[32];
  pdl_wait();
  const int b = blockIdx.x, tid = threadIdx.x;
  float* row = score + (int64_t)b * score_stride;

  float mx = -INFINITY;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float x = row[i] * scale;
    if (CACHE_ROW) row_cache[i] = x;
    mx = fmaxf(mx, x);
  }
  mx = block_reduce_max(mx, red);
  float sum = 0.f;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float x = CACHE_ROW ? row_cache[i] : row[i] * scale;
    const float e = expf(x - mx);
    if (CACHE_ROW) row_cache[i] = e;
    sum += e;
  }
  sum = block_reduce_sum(sum, red);
  const int kv_len = (cumulative != nullptr && kv_lens != nullptr) ? min(kv_lens[b], width) : 0;
  // cum_rows: row of batch entry b inside `cumulative` (batches need not occupy consecutive rows)
  float* cum = cumulative != nullptr ? cumulative + (int64_t)(cum_rows != nullptr ? cum_rows[b] : b) * cum_stride : nullptr;
  // refill != 0: the normalised row only feeds the cumulative update, so the step buffer is handed back
  // pre-filled for the next step's max-updates (sparse_controller.py:545-594 refills it with a launch of its own)
  const bool do_refill = refill != 0.f && cumulative != nullptr;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float e = CACHE_ROW ? row_cache[i] : expf(row[i] * scale - mx);
    const float pr = e / sum;
    row[i] = do_refill ? refill : pr;
    if (i < kv_len) cum[i] = (i < kv_len - 1) ? cum[i] + pr : pr;
  }
}

// fp32 bits -> 32-bit key that orders like the float value.
__device__ __forceinline__ uint32_t f32_order_key(uint32_t bits) {
  return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}

__device__ __forceinline__ int block_exclusive_scan_i(int v, int* warp_scan) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads();
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
    warp_scan[lane] = winc - w;
  }
  __syncthreads();
  return warp_scan[warp] + inc - v;
}

// One CTA per (layer,row) entry.  smem: keys[kv_len] u32, vals[kv_len] f32, slots[kv_len] i32.
__global__ void __launch_bounds__(kH2oThreads) h2o_select_compact_kernel(
    float* __restrict__ cumulative, const int64_t* __restrict__ cum_rows,
    int32_t* __restrict__ slot_tables, const int64_t* __restrict__ slot_rows,
    int32_t* __restrict__ free_stack, const int64_t* __restrict__ free_rows,
    int32_t* __restrict__ keep_out, const int32_t* __restrict__ kv_lens, int max_kv_len, int budget, int recent) {
  extern __shared__ __align__(16) unsigned char h2o_smem[];
  uint32_t* keys = reinterpret_cast<uint32_t*>(h2o_smem);
  float* vals = reinterpret_cast<float*>(keys + max_kv_len);
  int32_t* slots = reinterpret_cast<int32_t*>(vals + max_kv_len);
  __shared__ int hist[256];
  __shared__ int warp_scan[32];
  __shared__ int bcast[2];

  pdl_wait();
  const int r = blockIdx.x, tid = threadIdx.x;
  // rows of one burst may differ in length (the reference groups them by length, h2o.py:1498-1540)
  const int kv_len = kv_lens != nullptr ? min(kv_lens[r], max_kv_len) : max_kv_len;
  if (kv_len <= budget) return;
  float* cum = cumulative + cum_rows[r];
  int32_t* srow = slot_tables + slot_rows[r];
  const int recent_start = kv_len - recent;
  const int heavy = budget - recent;

  for (int i = tid; i < kv_len; i += kH2oThreads) {
    const float v = cum[i];
    vals[i] = v;
    keys[i] = f32_order_key(__float_as_uint(v));
    slots[i] = srow[i];
  }
  __syncthreads();

  // ---- radix select of the heavy-th largest key among [0, recent_start)
  uint32_t prefix = 0, prefix_mask = 0;
  int k = heavy;
  if (heavy > 0) {
    for (int shift = 24; shift >= 0; shift -= 8) {
      for (int i = tid; i < 256; i += kH2oThreads) hist[i] = 0;
      __syncthreads();
      for (int i = tid; i < recent_start; i += kH2oThreads) {
        const uint32_t key = keys[i];
        if ((key & prefix_mask) == prefix) atomicAdd(&hist[(key >> shift) & 0xff], 1);
      }
      __syncthreads();
      if (tid < 32) {
        const int lane = tid;
        int local[8], lsum = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          local[i] = hist[lane * 8 + i];
          lsum += local[i];
        }
        int suf = lsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int n = __shfl_down_sync(0xffffffffu, suf, o);
          if (lane + o < 32) suf += n;
        }
        const int above_lane = suf - lsum;
        if (above_lane < k && k <= suf) {
          int above = above_lane, bin = lane * 8 + 7;
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
      __syncthreads();
      prefix |= (uint32_t)bcast[0] << shift;
      prefix_mask |= 0xffu << shift;
      k -= bcast[1];
      __syncthreads();
    }
  }
  const uint32_t thresh = prefix;  // key of the heavy-th largest; k = ties to take (lowest index first)

  // ---- ordered compaction over contiguous chunks of [0, kv_len)
  const int per = (kv_len + kH2oThreads - 1) / kH2oThreads;
  const int i0 = min(tid * per, kv_len), i1 = min(i0 + per, kv_len);
  int cnt_eq = 0, cnt_sure = 0;
  for (int i = i0; i < i1; ++i) {
    if (i >= recent_start) {
      ++cnt_sure;
    } else if (heavy > 0) {
      cnt_sure += keys[i] > thresh;
      cnt_eq += keys[i] == thresh;
    }
  }
  const int eq_before = block_exclusive_scan_i(cnt_eq, warp_scan);
  const int eq_take = max(0, min(cnt_eq, k - eq_before));
  int pos = block_exclusive_scan_i(cnt_sure + eq_take, warp_scan);
  int drop = i0 - pos;  // dropped entries before this chunk
  int eq_left = eq_take;
  const int64_t free_base = free_rows != nullptr ? free_rows[r] : 0;
  for (int i = i0; i < i1; ++i) {
    bool keep = i >= recent_start;
    if (!keep && heavy > 0) {
      const uint32_t key = keys[i];
      keep = key > thresh;
      if (key == thresh && eq_left > 0) {
        keep = true;
        --eq_left;
      }
    }
    if (keep) {
      if (keep_out != nullptr) keep_out[(int64_t)r * budget + pos] = i;
      cum[pos] = vals[i];
      srow[pos] = slots[i];
      ++pos;
    } else {
      if (free_stack != nullptr) free_stack[free_base + drop] = slots[i];
      ++drop;
    }
  }
  __syncthreads();
  // zero the tails (snapkv.py:1797-1803 zeroes the slot tail; scores keep `budget` entries)
  for (int i = budget + tid; i < kv_len; i += kH2oThreads) {
    srow[i] = 0;
    cum[i] = 0.f;
  }
}

// =========================================================================
// Prefill window scoring (H2O / SnapKV-style token scores of a prefill chunk).
//
// Reference behaviour: kernels/triton/prefill_score.py:70-428 (probability mode: partial stats, reduce, final)
// driven by H2OCacheManager.collect_prefill_attention_score (engine/cache_manager/h2o.py:778-894):
//   for batch row b the queries q_abs in [score_start, score_end) (the last <= 128 tokens of the chunk) attend
//   causally to the row's kv positions in [candidate_start, max(candidate_start, ctx - num_recent));
//   score[b][kv] = max over q heads of ( sum over those queries of softmax_kv(q.k * D^-0.5)[q][kv] ) / n_queries.
// One CTA per (batch row, kv head, block of <= 16/tiles_per_head q heads): 8 warps x 2 row tiles of 16
// queries each (Q fragments live in registers), K tiles of 64 tokens staged through shared memory
// (slot-table indirect 16-byte loads), two passes over the row's K (softmax statistics, then probabilities),
// mma.sync m16n8k16 bf16 / fp32.  KV heads and head blocks meet in fp32 atomic max (scores are >= 0).
// =========================================================================
constexpr int kPsThreads = 256;
constexpr int kPsWarps = kPsThreads / 32;
constexpr int kPsTilesPerWarp = 2;
constexpr int kPsKTile = 64;                      // kv tokens per shared-memory tile
constexpr int kPsKStride = kRowBytes + 16;        // +16: ldmatrix rows hit distinct banks
constexpr float kPsMasked = -1.0e20f;

struct PrefillScoreArgs {
  const __nv_bfloat16* q;  // [T, Hq, 128], token stride q_stride elements
  int64_t q_stride;
  const __nv_bfloat16* k_cache;  // [slots, Hkv, 128]
  int64_t num_slots;
  const int32_t* slot_table;  // [rows][table_stride]
  int64_t table_stride;
  const int32_t* req_rows;      // [B] table row
  const int32_t* start_loc;     // [B] first chunk token of the row inside q
  const int32_t* context_lens;  // [B] physical row length including the chunk
  const int32_t* prompt_cache_lens;  // [B] physical row length before the chunk
  const int32_t* score_starts;  // [B] first scored query (absolute physical position)
  const int32_t* score_ends;    // [B]
  float* score;                 // [B][score_stride], zero-filled by the launcher
  int64_t score_stride;
  int hq, hkv, tiles_per_head, heads_per_cta, head_blocks;
  int candidate_start, num_recent;
  float sm_scale;
};

__global__ void __launch_bounds__(kPsThreads) h2o_prefill_score_kernel(PrefillScoreArgs a) {
  __shared__ __align__(16) unsigned char k_tile[kPsKTile * kPsKStride];
  __shared__ float head_acc[16][kPsKTile];  // per (head of this CTA, token of the tile): sum over queries of p
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int gsz = a.hq / a.hkv;
  const int hb = blockIdx.x % a.head_blocks;
  const int bkv = blockIdx.x / a.head_blocks;
  const int b = bkv / a.hkv, kvh = bkv % a.hkv;
  const int ctx = a.context_lens[b], cache_len = a.prompt_cache_lens[b];
  const int chunk_len = ctx - cache_len;
  const int q_start = a.score_starts[b], q_end = a.score_ends[b];
  const float inv_qlen = 1.f / (float)max(q_end - q_start, 1);
  const int cand_end = max(a.candidate_start, ctx - a.num_recent);
  const int32_t* table = a.slot_table + (int64_t)a.req_rows[b] * a.table_stride;
  const int64_t q_base = a.start_loc[b];

  // ---- this warp's row tiles: tile r -> (head of the CTA, 16 consecutive queries)
  uint32_t qa[kPsTilesPerWarp][8][4];
  int row_q[kPsTilesPerWarp][2];       // absolute query position of rows g, g + 8 (-1: invalid row)
  int tile_head[kPsTilesPerWarp];      // head index inside this CTA's block, -1: no such tile
#pragma unroll
  for (int x = 0; x < kPsTilesPerWarp; ++x) {
    const int r = warp + x * kPsWarps;
    const int hl = r / a.tiles_per_head, qt = r - hl * a.tiles_per_head;
    const int head_in_group = hb * a.heads_per_cta + hl;
    const bool tile_ok = hl < a.heads_per_cta && head_in_group < gsz;
    tile_head[x] = tile_ok ? hl : -1;
#pragma unroll
    for (int half = 0; half < 2; ++half) {
      const int q_abs = q_start + qt * 16 + g + half * 8;
      const int q_rel = q_abs - cache_len;
      const bool ok = tile_ok && q_abs < q_end && q_rel >= 0 && q_rel < chunk_len;
      row_q[x][half] = ok ? q_abs : -1;
      const uint32_t* qrow = reinterpret_cast<const uint32_t*>(
          a.q + (q_base + max(q_rel, 0)) * a.q_stride + (int64_t)(kvh * gsz + min(head_in_group, gsz - 1)) * kHeadDim);
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        qa[x][kk][half] = ok ? qrow[kk * 8 + t] : 0u;          // a0 / a1: dims 2t, 2t+1
        qa[x][kk][2 + half] = ok ? qrow[kk * 8 + 4 + t] : 0u;  // a2 / a3: dims 2t+8, 2t+9
      }
    }
  }

  float m_run[kPsTilesPerWarp][2], l_run[kPsTilesPerWarp][2];
#pragma unroll
  for (int x = 0; x < kPsTilesPerWarp; ++x) m_run[x][0] = m_run[x][1] = kPsMasked, l_run[x][0] = l_run[x][1] = 0.f;

  const int n_tiles = (min(cand_end, ctx) + kPsKTile - 1) / kPsKTile;
  for (int pass = 0; pass < 2; ++pass) {
    for (int kt = 0; kt < n_tiles; ++kt) {
      const int kv0 = kt * kPsKTile;
      // ---- stage 64 K rows of this kv head (16 x 16-byte segments each)
      __syncthreads();
      for (int c = tid; c < kPsKTile * 16; c += kPsThreads) {
        const int tok = c >> 4, seg = c & 15;
        const int kv = kv0 + tok;
        uint4 v = make_uint4(0, 0, 0, 0);
        if (kv >= a.candidate_start && kv < cand_end) {
          const int slot = table[kv];
          if (slot >= 0 && slot < a.num_slots)
            v = *reinterpret_cast<const uint4*>(a.k_cache + ((int64_t)slot * a.hkv + kvh) * kHeadDim + seg * 8);
        }
        *reinterpret_cast<uint4*>(k_tile + tok * kPsKStride + seg * 16) = v;
      }
      if (pass == 1)
        for (int c = tid; c < 16 * kPsKTile; c += kPsThreads) (&head_acc[0][0])[c] = 0.f;
      __syncthreads();
#pragma unroll
      for (int x = 0; x < kPsTilesPerWarp; ++x) {
        if (tile_head[x] < 0) continue;  // warp-uniform
        for (int sub = 0; sub < kPsKTile / 16; ++sub) {
          // S(16 queries x 16 tokens) = Q K^T: two n-tiles of 8 tokens
          float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
          const uint32_t kaddr = smem_u32(k_tile) + (sub * 16 + (lane & 7) + (lane >> 4) * 8) * kPsKStride + ((lane >> 3) & 1) * 16;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            uint32_t kb[4];  // {b0, b1} of tokens 0-7, {b0, b1} of tokens 8-15
            ldmatrix_x4(kb, kaddr + kk * 32);
            mma_16816(sc[0], qa[x][kk], kb[0], kb[1]);
            mma_16816(sc[1], qa[x][kk], kb[2], kb[3]);
          }
          // c0:(row g, tok 2t) c1:(g, 2t+1) c2:(g+8, 2t) c3:(g+8, 2t+1)
          float p[2][4];
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const int kv = kv0 + sub * 16 + nt * 8 + 2 * t + (e & 1);
              const int qpos = row_q[x][e >> 1];
              const bool valid = qpos >= 0 && kv >= a.candidate_start && kv < cand_end && qpos >= kv;
              p[nt][e] = valid ? sc[nt][e] * a.sm_scale : kPsMasked;
            }
          if (pass == 0) {
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              float mx = fmaxf(fmaxf(p[0][2 * half], p[0][2 * half + 1]), fmaxf(p[1][2 * half], p[1][2 * half + 1]));
              mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
              mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
              const float mnew = fmaxf(m_run[x][half], mx);
              float sum = 0.f;
#pragma unroll
              for (int nt = 0; nt < 2; ++nt)
#pragma unroll
                for (int e = 0; e < 2; ++e) {
                  const float v = p[nt][2 * half + e];
                  sum += v > 0.5f * kPsMasked ? expf(v - mnew) : 0.f;
                }
              sum += __shfl_xor_sync(0xffffffffu, sum, 1);
              sum += __shfl_xor_sync(0xffffffffu, sum, 2);
              l_run[x][half] = l_run[x][half] * expf(m_run[x][half] - mnew) + sum;
              m_run[x][half] = mnew;
            }
          } else {
            // probabilities, summed over the tile's 16 queries per token, added to the head's accumulator
#pragma unroll
            for (int nt = 0; nt < 2; ++nt)
#pragma unroll
              for (int e = 0; e < 2; ++e) {
                float col = 0.f;
#pragma unroll
                for (int half = 0; half < 2; ++half) {
                  const float v = p[nt][2 * half + e];
                  const float l = l_run[x][half];
                  col += v > 0.5f * kPsMasked ? expf(v - m_run[x][half]) / (l > 0.f ? l : 1.f) : 0.f;
                }
                col += __shfl_xor_sync(0xffffffffu, col, 4);
                col += __shfl_xor_sync(0xffffffffu, col, 8);
                col += __shfl_xor_sync(0xffffffffu, col, 16);
                if (g == 0 && col != 0.f) atomicAdd(&head_acc[tile_head[x]][sub * 16 + nt * 8 + 2 * t + e], col);
              }
          }
        }
      }
      if (pass == 1) {
        __syncthreads();
        if (tid < kPsKTile) {
          const int kv = kv0 + tid;
          if (kv >= a.candidate_start && kv < cand_end) {
            float best = 0.f;
            for (int h = 0; h < a.heads_per_cta; ++h) best = fmaxf(best, head_acc[h][tid] * inv_qlen);
            if (best > 0.f) atomicMax(reinterpret_cast<int*>(a.score + (int64_t)b * a.score_stride + kv), __float_as_int(best));
          }
        }
      }
    }
  }
}

// cumulative[row][i] = (i < prev_len ? cumulative[row][i] : 0) + weight * score[b][i] for i < new_len
// (H2OCacheManager._accumulate_score, h2o.py:611-626, with _expand_score's zero extension).
__global__ void __launch_bounds__(256) h2o_prefill_accumulate_kernel(
    const float* __restrict__ score, int64_t score_stride, float* __restrict__ cumulative, int64_t cum_stride,
    const int32_t* __restrict__ cum_rows, const int32_t* __restrict__ prev_lens, const int32_t* __restrict__ new_lens,
    const int32_t* __restrict__ score_starts, const int32_t* __restrict__ score_ends) {
  pdl_wait();
  const int b = blockIdx.x;
  const int prev = prev_lens[b], n = new_lens[b];
  const float weight = (float)(score_ends[b] - score_starts[b]);
  const float* srow = score + (int64_t)b * score_stride;
  float* cum = cumulative + (int64_t)cum_rows[b] * cum_stride;
  for (int i = threadIdx.x; i < n; i += blockDim.x) cum[i] = (i < prev ? cum[i] : 0.f) + weight * srow[i];
}

int launch_h2o_prefill_score(const PrefillScoreHostArgs& h, cudaStream_t stream) {
  if (h.batch <= 0 || h.max_query_len <= 0 || h.max_context_len <= 0) return 0;
  PrefillScoreArgs a{};
  a.q = (const __nv_bfloat16*)h.q; a.q_stride = h.q_stride; a.k_cache = (const __nv_bfloat16*)h.k_cache;
  a.num_slots = h.num_slots; a.slot_table = h.slot_table; a.table_stride = h.table_stride; a.req_rows = h.req_rows;
  a.start_loc = h.start_loc; a.context_lens = h.context_lens; a.prompt_cache_lens = h.prompt_cache_lens;
  a.score_starts = h.score_starts; a.score_ends = h.score_ends; a.score = h.score; a.score_stride = h.score_stride;
  a.hq = h.hq; a.hkv = h.hkv; a.candidate_start = h.candidate_start; a.num_recent = h.num_recent;
  a.sm_scale = h.sm_scale;
  a.tiles_per_head = (h.max_query_len + 15) / 16;
  const int gsz = h.hq / h.hkv;
  a.heads_per_cta = kPsWarps * kPsTilesPerWarp / a.tiles_per_head;
  if (a.heads_per_cta > gsz) a.heads_per_cta = gsz;
  a.head_blocks = (gsz + a.heads_per_cta - 1) / a.heads_per_cta;
  cudaError_t e = cudaMemsetAsync(h.score, 0, sizeof(float) * (size_t)h.score_stride * h.batch, stream);
  if (e != cudaSuccess) return cuda_error("h2o_prefill_score memset", e);
  e = launch_kernel(h2o_prefill_score_kernel, dim3(h.batch * h.hkv * a.head_blocks), dim3(kPsThreads), 0, stream, a);
  if (e != cudaSuccess) return cuda_error("h2o_prefill_score launch", e);
  if (h.cumulative != nullptr) {
    e = launch_kernel(h2o_prefill_accumulate_kernel, dim3(h.batch), dim3(256), 0, stream, (const float*)h.score,
                      h.score_stride, h.cumulative, h.cum_stride, h.cum_rows, h.prompt_cache_lens, h.context_lens,
                      h.score_starts, h.score_ends);
    if (e != cudaSuccess) return cuda_error("h2o_prefill_accumulate launch", e);
  }
  return check_launch("h2o_prefill_score");
}

// =========================================================================
// Host -> device hand-over of one H2O decode step (H2OCacheManager.prepare_decode_static, h2o.py:257-476):
// meta = [rows[B], new_lens[B], free_top] packed by the host (one pinned copy); thread (l, b) pops the b-th of
// the B topmost free slots of layer l, appends it to the layer's slot row and fills the static step buffers.
// =========================================================================
__global__ void __launch_bounds__(256) h2o_prepare_decode_kernel(
    const int32_t* __restrict__ meta, const int32_t* __restrict__ free_stack, int64_t free_stride,
    int32_t* __restrict__ slot_tables, int64_t layer_stride, int64_t row_stride,
    int32_t* __restrict__ layer_slot_mapping, int64_t lsm_stride, int32_t* __restrict__ context_lens,
    int32_t* __restrict__ req_indices, int32_t* __restrict__ slot_mapping, int32_t* __restrict__ kv_lens,
    int32_t* __restrict__ rows_out, int batch, int layers) {
  pdl_wait();
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= batch * layers) return;
  const int l = idx / batch, b = idx - l * batch;
  const int row = meta[b], len = meta[batch + b], free_top = meta[2 * batch];
  const int32_t slot = free_stack[(int64_t)l * free_stride + free_top - batch + b];
  slot_tables[(int64_t)l * layer_stride + (int64_t)row * row_stride + (len - 1)] = slot;
  layer_slot_mapping[(int64_t)l * lsm_stride + b] = slot;
  if (l == 0) {
    context_lens[b] = len;
    req_indices[b] = row;
    slot_mapping[b] = slot;
    kv_lens[b] = len;
    rows_out[b] = row;
  }
}

int launch_h2o_prepare_decode(const int32_t* meta, const int32_t* free_stack, int64_t free_stride, int32_t* slot_tables,
                              int64_t layer_stride, int64_t row_stride, int32_t* layer_slot_mapping, int64_t lsm_stride,
                              int32_t* context_lens, int32_t* req_indices, int32_t* slot_mapping, int32_t* kv_lens,
                              int32_t* rows_out, int batch, int layers, cudaStream_t stream) {
  if (batch <= 0 || layers <= 0) return 0;
  const int n = batch * layers;
  const cudaError_t e = launch_kernel(h2o_prepare_decode_kernel, dim3((n + 255) / 256), dim3(256), 0, stream, meta,
                                      free_stack, free_stride, slot_tables, layer_stride, row_stride, layer_slot_mapping,
                                      lsm_stride, context_lens, req_indices, slot_mapping, kv_lens, rows_out, batch, layers);
  if (e != cudaSuccess) return cuda_error("h2o_prepare_decode launch", e);
  return check_launch("h2o_prepare_decode");
}

int launch_h2o_normalize_accumulate(float* score, int64_t score_stride, int batch, int width,
                                    float scale, float* cumulative, int64_t cum_stride,
                                    const int32_t* cum_rows, const int32_t* kv_lens, float refill,
                                    cudaStream_t stream) {
  if (batch <= 0 || width <= 0) return 0;
  const size_t smem = (size_t)width * sizeof(float);
  if (smem <= 96 * 1024) {
    auto kern = h2o_normalize_accumulate_kernel<true>;
    if (smem > 32 * 1024) {  // static + dynamic shared memory share the 48 KB default limit
      cudaError_t e =
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return cuda_error("h2o_normalize smem attr", e);
    }
    const cudaError_t e = launch_kernel(kern, dim3(batch), dim3(kH2oThreads), smem, stream, score, score_stride, width,
                                        scale, cumulative, cum_stride, cum_rows, kv_lens, refill);
    if (e != cudaSuccess) return cuda_error("h2o_normalize_accumulate launch", e);
  } else {
    const cudaError_t e = launch_kernel(h2o_normalize_accumulate_kernel<false>, dim3(batch), dim3(kH2oThreads), 0, stream,
                                        score, score_stride, width, scale, cumulative, cum_stride, cum_rows, kv_lens,
                                        refill);
    if (e != cudaSuccess) return cuda_error("h2o_normalize_accumulate launch", e);
  }
  return check_launch("h2o_normalize_accumulate");
}

int launch_h2o_select_compact(float* cumulative, const int64_t* cum_rows, int32_t* slot_tables,
                              const int64_t* slot_rows, int32_t* free_stack,
                              const int64_t* free_rows, int32_t* keep_out, int num_rows,
                              const int32_t* kv_lens, int kv_len, int budget, int recent, cudaStream_t stream) {
  if (num_rows <= 0) return 0;
  const size_t smem = (size_t)kv_len * 12;
  auto kern = h2o_select_compact_kernel;
  if (smem > 32 * 1024) {  // static + dynamic shared memory share the 48 KB default limit
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_error("h2o_select_compact smem attr", e);
  }
  const cudaError_t e = launch_kernel(kern, dim3(num_rows), dim3(kH2oThreads), smem, stream, cumulative, cum_rows,
                                      slot_tables, slot_rows, free_stack, free_rows, keep_out, kv_lens, kv_len, budget,
                                      recent);
  if (e != cudaSuccess) return cuda_error("h2o_select_compact launch", e);
  return check_launch("h2o_select_compact");
}

}  // namespace svb200
