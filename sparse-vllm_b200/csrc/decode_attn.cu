// Sparse paged GQA decode attention for sm_100a: split-K over the visible token
// slots of each row, slot-table-indirect K/V staging, tensor-core QK^T / PV tile
// contractions, fp32 online softmax, optional H2O head-max raw-score epilogue and a
// fused split-K combine (the last CTA of a row merges the partials).
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/kernels/triton/):
//   gqa_flash_decoding_stage1.py:6-121 (stage 1), 211-295 (2D head-max score),
//   flash_decoding_stage2.py:8-46 (combine).  Conventions kept: logits scaled by
//   head_dim^-0.5, softmax numerators cast to bf16 before PV, per-split partials
//   normalised by their own sum, lse = max + log(sum), empty split -> (0, -inf).
//
// Work decomposition: an item is (batch row b, split s, head group hg of HPC KV
// heads).  A persistent CTA (8 warps) walks items round-robin; the K/V tiles of
// consecutive items form one stream so the next tile is always in flight while the
// current one is consumed.  Inside a tile each warp owns one KV head and one
// 16-token slice: S^T(16 tok x 8 heads) = K Q^T, then O^T(128 x 8) += V^T P^T.
#include "common.cuh"
#include "launch.h"

namespace svb200 {

constexpr int kAttnWarps = 8;
constexpr int kAttnThreads = kAttnWarps * 32;
constexpr int kAttnStages = 2;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;

template <int HPC>
struct AttnGeom {
  static constexpr int TL = kAttnWarps / HPC;       // warps (token lanes) per KV head
  static constexpr int TILE_TOK = 16 * TL;          // tokens per staged tile
  static constexpr int TOK_BYTES = HPC * kRowBytes;  // bytes of one token's head-group row
  static constexpr int TOK_STRIDE = TOK_BYTES + 16;  // padded: consecutive tokens shift 4 banks
  static constexpr int TILE_BYTES = TILE_TOK * TOK_STRIDE;
  static constexpr int CHUNKS_PER_TOK = HPC * 16;   // 16-byte chunks per token row
  static constexpr int CHUNKS = TILE_TOK * CHUNKS_PER_TOK;  // = 2048
  static_assert(CHUNKS % kAttnThreads == 0, "tile must divide over the CTA");
};

// Position of a tile in this CTA's stream.
struct TileCursor {
  int item;    // global item id; >= n_items means end of stream
  int tile;    // tile inside the item
  int ntiles;  // tiles of the item
  int b, hg, split;
  int t0, t1;  // token range of the item
  int64_t row;
};

struct AttnDims {
  int n_items, head_groups, num_splits, split_tokens;
};

template <int HPC>
__device__ __forceinline__ void cursor_load_item(TileCursor& c, const AttnDims& d,
                                                 const int32_t* __restrict__ req_rows,
                                                 const int32_t* __restrict__ seq_lens) {
  // item = (b * num_splits + split) * head_groups + hg
  c.hg = c.item % d.head_groups;
  const int bs = c.item / d.head_groups;
  c.split = bs % d.num_splits;
  c.b = bs / d.num_splits;
  const int len = seq_lens[c.b];
  c.row = req_rows[c.b];
  c.t0 = c.split * d.split_tokens;
  c.t1 = min(len, c.t0 + d.split_tokens);
  const int n = c.t1 - c.t0;
  c.ntiles = n > 0 ? (n + AttnGeom<HPC>::TILE_TOK - 1) / AttnGeom<HPC>::TILE_TOK : 0;
  c.tile = 0;
}

// Advance to the next tile of the stream (skipping items without tiles).
template <int HPC>
__device__ __forceinline__ void cursor_next(TileCursor& c, const AttnDims& d,
                                            const int32_t* __restrict__ req_rows,
                                            const int32_t* __restrict__ seq_lens) {
  if (c.item >= d.n_items) return;
  if (++c.tile < c.ntiles) return;
  for (;;) {
    c.item += gridDim.x;
    if (c.item >= d.n_items) {
      c.ntiles = 0;
      c.tile = 0;
      return;
    }
    cursor_load_item<HPC>(c, d, req_rows, seq_lens);
    if (c.ntiles > 0) return;
  }
}

template <int HPC>
__device__ __forceinline__ void cursor_first(TileCursor& c, const AttnDims& d,
                                             const int32_t* __restrict__ req_rows,
                                             const int32_t* __restrict__ seq_lens) {
  c.item = blockIdx.x;
  c.tile = 0;
  c.ntiles = 0;
  while (c.item < d.n_items) {
    cursor_load_item<HPC>(c, d, req_rows, seq_lens);
    if (c.ntiles > 0) return;
    c.item += gridDim.x;
  }
}

// Slot of token `tok` of the cursor's tile, or -1 when outside the item / invalid.
template <int HPC>
__device__ __forceinline__ int cursor_slot(const TileCursor& c, int n_items, int tok,
                                           const int32_t* __restrict__ slot_table,
                                           int64_t slot_stride, int64_t num_slots) {
  if (c.item >= n_items) return -1;
  const int t = c.t0 + c.tile * AttnGeom<HPC>::TILE_TOK + tok;
  if (t >= c.t1) return -1;
  const int s = slot_table[c.row * slot_stride + t];
  return (s >= 0 && s < num_slots) ? s : -1;
}

template <int HPC, int NT, bool WITH_SCORE>
__global__ void __launch_bounds__(kAttnThreads, 1) sparse_decode_attn_kernel(AttnParams p) {
  using G = AttnGeom<HPC>;
  constexpr int TL = G::TL;

  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* sK = smem;                                   // [stages][TILE_BYTES]
  unsigned char* sV = sK + kAttnStages * G::TILE_BYTES;        // [stages][TILE_BYTES]
  int* sSlot = reinterpret_cast<int*>(sV + kAttnStages * G::TILE_BYTES);  // [3][TILE_TOK]
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(sSlot + 3 * G::TILE_TOK);  // [HPC*gsz][128]
  const int gsz = p.hq / p.hkv;
  float* sMerge = reinterpret_cast<float*>(sQ + HPC * gsz * kHeadDim);  // [NREG][256] (TL > 1)
  __shared__ int s_flag;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int h_local = warp % HPC, tl = warp / HPC;
  const int tok0 = tl * 16;

  AttnDims dims;
  dims.head_groups = p.hkv / HPC;
  dims.num_splits = p.num_splits;
  dims.split_tokens = p.split_tokens;
  dims.n_items = p.batch * p.num_splits * dims.head_groups;

  const __nv_bfloat16* kc = reinterpret_cast<const __nv_bfloat16*>(p.k_cache);
  const __nv_bfloat16* vc = reinterpret_cast<const __nv_bfloat16*>(p.v_cache);
  const int64_t slot_elems = (int64_t)p.hkv * kHeadDim;
  const float scale_log2 = p.sm_scale * kLog2e;

  // Issue the cp.async copies of one tile (slots already staged in sSlot[ring]).
  auto issue_tile = [&](const TileCursor& c, int stage, int ring) {
    if (c.item < dims.n_items) {
      const uint32_t kdst = smem_u32(sK + stage * G::TILE_BYTES);
      const uint32_t vdst = smem_u32(sV + stage * G::TILE_BYTES);
      const int64_t head_off = (int64_t)c.hg * HPC * kHeadDim;
#pragma unroll
      for (int i = 0; i < G::CHUNKS / kAttnThreads; ++i) {
        const int ch = i * kAttnThreads + tid;
        const int tok = ch / G::CHUNKS_PER_TOK, off = ch % G::CHUNKS_PER_TOK;
        const int slot = sSlot[ring * G::TILE_TOK + tok];
        const int64_t src = (int64_t)max(slot, 0) * slot_elems + head_off + off * 8;
        const uint32_t d = tok * G::TOK_STRIDE + off * 16;
        const int nbytes = slot >= 0 ? 16 : 0;  // invalid token -> zero fill
        cp_async_16(kdst + d, kc + src, nbytes);
        cp_async_16(vdst + d, vc + src, nbytes);
      }
    }
    cp_async_commit();
  };

  TileCursor c0, c1, c2;
  cursor_first<HPC>(c0, dims, p.req_rows, p.seq_lens);
  c1 = c0;
  cursor_next<HPC>(c1, dims, p.req_rows, p.seq_lens);
  c2 = c1;
  cursor_next<HPC>(c2, dims, p.req_rows, p.seq_lens);

  // ---- stream prologue: slots of tiles 0 and 1, data of tile 0
  if (tid < G::TILE_TOK) {
    sSlot[0 * G::TILE_TOK + tid] =
        cursor_slot<HPC>(c0, dims.n_items, tid, p.slot_table, p.slot_table_stride, p.num_slots);
    sSlot[1 * G::TILE_TOK + tid] =
        cursor_slot<HPC>(c1, dims.n_items, tid, p.slot_table, p.slot_table_stride, p.num_slots);
  }
  __syncthreads();
  issue_tile(c0, 0, 0);

  int gtile = 0;  // index of the current tile in this CTA's stream

  // ---- item loop (items without tiles still publish an empty partial)
  for (int item = blockIdx.x; item < dims.n_items; item += gridDim.x) {
    const int hg = item % dims.head_groups;
    const int bs = item / dims.head_groups;
    const int split = bs % dims.num_splits;
    const int b = bs / dims.num_splits;
    const int len = p.seq_lens[b];
    const int it0 = split * dims.split_tokens;
    const int it1 = min(len, it0 + dims.split_tokens);
    const int ntiles = it1 > it0 ? (it1 - it0 + G::TILE_TOK - 1) / G::TILE_TOK : 0;

    // q rows of this head group -> smem -> B fragments
    uint32_t qb[NT][8][2];
    if (ntiles > 0) {
      const uint4* qsrc = reinterpret_cast<const uint4*>(
          reinterpret_cast<const __nv_bfloat16*>(p.q) + (int64_t)b * p.q_stride_b +
          (int64_t)hg * HPC * gsz * kHeadDim);
      uint4* qdst = reinterpret_cast<uint4*>(sQ);
      for (int i = tid; i < HPC * gsz * 16; i += kAttnThreads) qdst[i] = qsrc[i];
      __syncthreads();
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        const int hi = nt * 8 + g;
        const uint32_t* qrow =
            reinterpret_cast<const uint32_t*>(sQ + (h_local * gsz + min(hi, gsz - 1)) * kHeadDim);
#pragma unroll
        for (int kk = 0; kk < 8; ++kk) {
          qb[nt][kk][0] = hi < gsz ? qrow[kk * 8 + t] : 0u;
          qb[nt][kk][1] = hi < gsz ? qrow[kk * 8 + 4 + t] : 0u;
        }
      }
    }

    float m_run[NT][2], l_run[NT][2], o_acc[NT][8][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      m_run[nt][0] = m_run[nt][1] = -INFINITY;
      l_run[nt][0] = l_run[nt][1] = 0.f;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt)
#pragma unroll
        for (int i = 0; i < 4; ++i) o_acc[nt][dt][i] = 0.f;
    }

    for (int tile = 0; tile < ntiles; ++tile, ++gtile) {
      const int stage = gtile & 1, ring = gtile % 3;
      // slots of the tile after next (consumed at the end of this iteration)
      int pre_slot = -1;
      if (tid < G::TILE_TOK)
        pre_slot =
            cursor_slot<HPC>(c2, dims.n_items, tid, p.slot_table, p.slot_table_stride, p.num_slots);
      // next tile's data into the other stage
      issue_tile(c1, stage ^ 1, (gtile + 1) % 3);
      cp_async_wait<1>();
      __syncthreads();

      // ------------------------------------------------ compute on `stage`
      const int sl0 = sSlot[ring * G::TILE_TOK + tok0 + g];
      const int sl1 = sSlot[ring * G::TILE_TOK + tok0 + g + 8];
      const bool tv0 = sl0 >= 0, tv1 = sl1 >= 0;
      if (__any_sync(0xffffffffu, tv0 || tv1)) {
        const uint32_t kbase = smem_u32(sK + stage * G::TILE_BYTES) + h_local * kRowBytes;
        const uint32_t vbase = smem_u32(sV + stage * G::TILE_BYTES) + h_local * kRowBytes;
        float s[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
        {
          const uint32_t kaddr =
              kbase + (tok0 + (lane & 7) + ((lane >> 3) & 1) * 8) * G::TOK_STRIDE + (lane >> 4) * 16;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            uint32_t a[4];
            ldmatrix_x4(a, kaddr + kk * 32);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816(s[nt], a, qb[nt][kk][0], qb[nt][kk][1]);
          }
        }
        if constexpr (WITH_SCORE) {
          // raw (unscaled) max over the heads of this KV head, then over KV heads via atomics
          float r0 = -INFINITY, r1 = -INFINITY;
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            if (col < gsz) r0 = fmaxf(r0, s[nt][0]), r1 = fmaxf(r1, s[nt][2]);
            if (col + 1 < gsz) r0 = fmaxf(r0, s[nt][1]), r1 = fmaxf(r1, s[nt][3]);
          }
          r0 = fmaxf(r0, __shfl_xor_sync(0xffffffffu, r0, 1));
          r0 = fmaxf(r0, __shfl_xor_sync(0xffffffffu, r0, 2));
          r1 = fmaxf(r1, __shfl_xor_sync(0xffffffffu, r1, 1));
          r1 = fmaxf(r1, __shfl_xor_sync(0xffffffffu, r1, 2));
          if (t == 0) {
            float* srow = p.attn_score + (int64_t)b * p.score_stride + it0 + tile * G::TILE_TOK + tok0;
            if (tv0) atomic_max_f32(srow + g, r0);
            if (tv1) atomic_max_f32(srow + g + 8, r1);
          }
        }
        uint32_t pb[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float x[4];
          x[0] = tv0 ? s[nt][0] * scale_log2 : -INFINITY;
          x[1] = tv0 ? s[nt][1] * scale_log2 : -INFINITY;
          x[2] = tv1 ? s[nt][2] * scale_log2 : -INFINITY;
          x[3] = tv1 ? s[nt][3] * scale_log2 : -INFINITY;
          float alpha[2];
#pragma unroll
          for (int j = 0; j < 2; ++j) {
            float mloc = fmaxf(x[j], x[j + 2]);
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 4));
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 8));
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 16));
            const float mnew = fmaxf(m_run[nt][j], mloc);
            const float msafe = mnew == -INFINITY ? 0.f : mnew;
            alpha[j] = fast_exp2(m_run[nt][j] - msafe);
            x[j] = fast_exp2(x[j] - msafe);
            x[j + 2] = fast_exp2(x[j + 2] - msafe);
            l_run[nt][j] = l_run[nt][j] * alpha[j] + x[j] + x[j + 2];
            m_run[nt][j] = mnew;
          }
#pragma unroll
          for (int dt = 0; dt < 8; ++dt) {
            o_acc[nt][dt][0] *= alpha[0];
            o_acc[nt][dt][1] *= alpha[1];
            o_acc[nt][dt][2] *= alpha[0];
            o_acc[nt][dt][3] *= alpha[1];
          }
          // P (tok x head) -> bf16 -> transpose -> B fragment (tok as k, head as n)
          pb[nt][0] = movmatrix_trans(pack_bf16x2(x[0], x[1]));
          pb[nt][1] = movmatrix_trans(pack_bf16x2(x[2], x[3]));
        }
        {
          const uint32_t vaddr =
              vbase + (tok0 + (lane & 7) + (lane >> 4) * 8) * G::TOK_STRIDE + ((lane >> 3) & 1) * 16;
#pragma unroll
          for (int dt = 0; dt < 8; ++dt) {
            uint32_t a[4];
            ldmatrix_x4_trans(a, vaddr + dt * 32);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816(o_acc[nt][dt], a, pb[nt][0], pb[nt][1]);
          }
        }
      }
      // ------------------------------------------------ rotate the stream
      if (tid < G::TILE_TOK) sSlot[((gtile + 2) % 3) * G::TILE_TOK + tid] = pre_slot;
      c1 = c2;
      cursor_next<HPC>(c2, dims, p.req_rows, p.seq_lens);
      __syncthreads();
    }

    // ------------------------------------------------------ item epilogue
    // per-head sums live as per-lane partials: reduce over the 8 token lanes
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        float l = l_run[nt][j];
        l += __shfl_xor_sync(0xffffffffu, l, 4);
        l += __shfl_xor_sync(0xffffffffu, l, 8);
        l += __shfl_xor_sync(0xffffffffu, l, 16);
        l_run[nt][j] = l;
      }
    if constexpr (TL > 1) {
      if (ntiles > 0) {
        if (tl > 0) {
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            float* rec = sMerge + (nt * 36) * kAttnThreads + tid;
            rec[0 * kAttnThreads] = m_run[nt][0];
            rec[1 * kAttnThreads] = m_run[nt][1];
            rec[2 * kAttnThreads] = l_run[nt][0];
            rec[3 * kAttnThreads] = l_run[nt][1];
#pragma unroll
            for (int dt = 0; dt < 8; ++dt)
#pragma unroll
              for (int i = 0; i < 4; ++i) rec[(4 + dt * 4 + i) * kAttnThreads] = o_acc[nt][dt][i];
          }
        }
        __syncthreads();
        if (tl == 0) {
          for (int o = 1; o < TL; ++o) {
            const int ptid = (o * HPC + h_local) * 32 + lane;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
              const float* rec = sMerge + (nt * 36) * kAttnThreads + ptid;
              float sa[2], sb[2];
#pragma unroll
              for (int j = 0; j < 2; ++j) {
                const float mo = rec[j * kAttnThreads], lo = rec[(2 + j) * kAttnThreads];
                const float mnew = fmaxf(m_run[nt][j], mo);
                const float msafe = mnew == -INFINITY ? 0.f : mnew;
                sa[j] = fast_exp2(m_run[nt][j] - msafe);
                sb[j] = fast_exp2(mo - msafe);
                l_run[nt][j] = l_run[nt][j] * sa[j] + lo * sb[j];
                m_run[nt][j] = mnew;
              }
#pragma unroll
              for (int dt = 0; dt < 8; ++dt)
#pragma unroll
                for (int i = 0; i < 4; ++i)
                  o_acc[nt][dt][i] = o_acc[nt][dt][i] * sa[i & 1] +
                                     rec[(4 + dt * 4 + i) * kAttnThreads] * sb[i & 1];
            }
          }
        }
      }
    }
    // publish the partial of this (b, split, head group): normalised o, natural-log lse
    if (tl == 0) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int hi = nt * 8 + 2 * t + j;
          if (hi < gsz) {
            const int head = (hg * HPC + h_local) * gsz + hi;
            const float l = l_run[nt][j];
            const float inv = l > 0.f ? 1.f / l : 0.f;
            if (p.num_splits == 1) {
              __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                    (int64_t)b * p.out_stride_b + (int64_t)head * kHeadDim;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = __float2bfloat16_rn(o_acc[nt][dt][j] * inv);
                orow[dt * 16 + g + 8] = __float2bfloat16_rn(o_acc[nt][dt][2 + j] * inv);
              }
            } else {
              float* orow = p.mid_o + (int64_t)b * p.mo_sb + (int64_t)head * p.mo_sh +
                            (int64_t)split * p.mo_ss;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = o_acc[nt][dt][j] * inv;
                orow[dt * 16 + g + 8] = o_acc[nt][dt][2 + j] * inv;
              }
              if (g == 0)
                p.mid_lse[(int64_t)b * p.ml_sb + (int64_t)head * p.ml_sh + split] =
                    l > 0.f ? m_run[nt][j] * kLn2 + __logf(l) : -INFINITY;
            }
          }
        }
    }
    if (p.num_splits > 1) {
      // arrival ticket; the last split of (b, hg) merges all partials
      __threadfence();
      __syncthreads();
      if (tid == 0) {
        int* tk = p.tickets + b * dims.head_groups + hg;
        const int ticket = atomicAdd(tk, 1);
        s_flag = (ticket == p.num_splits - 1);
        if (s_flag) *tk = 0;
      }
      __syncthreads();
      if (s_flag) {
        __threadfence();
        const int nheads = HPC * gsz;
        for (int w = tid; w < nheads * 32; w += kAttnThreads) {
          const int head = hg * nheads + (w >> 5), d4 = w & 31;
          const float* lse_row = p.mid_lse + (int64_t)b * p.ml_sb + (int64_t)head * p.ml_sh;
          const float* o_row = p.mid_o + (int64_t)b * p.mo_sb + (int64_t)head * p.mo_sh;
          float mx = -INFINITY;
          for (int sidx = 0; sidx < p.num_splits; ++sidx)
            mx = fmaxf(mx, __ldcg(lse_row + sidx));
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          float wsum = 0.f;
          for (int sidx = 0; sidx < p.num_splits; ++sidx) {
            const float lse = __ldcg(lse_row + sidx);
            if (lse == -INFINITY) continue;
            const float wgt = __expf(lse - mx);
            const float4 v =
                __ldcg(reinterpret_cast<const float4*>(o_row + (int64_t)sidx * p.mo_ss) + d4);
            acc.x += wgt * v.x, acc.y += wgt * v.y, acc.z += wgt * v.z, acc.w += wgt * v.w;
            wsum += wgt;
          }
          const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
          uint2 o2;
          o2.x = pack_bf16x2(acc.x * inv, acc.y * inv);
          o2.y = pack_bf16x2(acc.z * inv, acc.w * inv);
          __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                (int64_t)b * p.out_stride_b + (int64_t)head * kHeadDim;
          *reinterpret_cast<uint2*>(orow + d4 * 4) = o2;
        }
      }
    }
    __syncthreads();  // sQ / sMerge / s_flag are reused by the next item
  }
  cp_async_wait<0>();
}

template <int HPC, int NT, bool WITH_SCORE>
static int launch_variant(const AttnParams& p, cudaStream_t stream) {
  using G = AttnGeom<HPC>;
  const int gsz = p.hq / p.hkv;
  size_t smem = (size_t)2 * kAttnStages * G::TILE_BYTES + 3 * G::TILE_TOK * sizeof(int) +
                (size_t)HPC * gsz * kRowBytes;
  if (G::TL > 1) smem += (size_t)NT * 36 * kAttnThreads * sizeof(float);
  auto kern = sparse_decode_attn_kernel<HPC, NT, WITH_SCORE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_error("sparse_decode_attn smem attr", e);
  const int n_items = p.batch * p.num_splits * (p.hkv / HPC);
  int grid = sm_count();
  if (grid > n_items) grid = n_items;
  kern<<<grid, kAttnThreads, smem, stream>>>(p);
  return check_launch("sparse_decode_attn");
}

template <int HPC>
static int launch_hpc(const AttnParams& p, cudaStream_t stream) {
  const int gsz = p.hq / p.hkv;
  const bool score = p.attn_score != nullptr;
  if (gsz <= 8)
    return score ? launch_variant<HPC, 1, true>(p, stream) : launch_variant<HPC, 1, false>(p, stream);
  return score ? launch_variant<HPC, 2, true>(p, stream) : launch_variant<HPC, 2, false>(p, stream);
}

int launch_sparse_decode_attn(const AttnParams& p, int heads_per_cta, cudaStream_t stream) {
  if (p.batch <= 0) return 0;
  int hpc = heads_per_cta;
  if (hpc <= 0) {
    // auto: whole rows of contiguous bytes per token as long as there are enough items
    hpc = 1;
    for (int c = 8; c >= 1; c >>= 1) {
      if (p.hkv % c) continue;
      const int items = p.batch * p.num_splits * (p.hkv / c);
      if (items >= sm_count() || c == 1) {
        hpc = c;
        break;
      }
    }
  }
  if (p.hkv % hpc) {
    set_error("sparse_decode_attn: heads_per_cta=%d does not divide num_kv_heads=%d", hpc, p.hkv);
    return 1;
  }
  switch (hpc) {
    case 1: return launch_hpc<1>(p, stream);
    case 2: return launch_hpc<2>(p, stream);
    case 4: return launch_hpc<4>(p, stream);
    case 8: return launch_hpc<8>(p, stream);
  }
  set_error("sparse_decode_attn: heads_per_cta must be 1, 2, 4 or 8 (got %d)", hpc);
  return 1;
}

}  // namespace svb200
