// Dense causal prefill attention over a paged KV row (chunked prefill with a cached prefix) for sm_100a.
//
// Reference behaviour (CURRENTF/Sparse-vLLM): kernels/triton/context_flashattention_nopad.py:11-79 (_fwd_kernel)
// and its wrapper 242-302: for batch row b the chunk's queries m in [0, chunk_len) sit at physical positions
// m + prompt_cache_len and attend to the row's kv positions n <= m + prompt_cache_len through the slot table
// (the chunk's own K/V rows are already stored); logits q.k * D^-0.5, online softmax in fp32, P cast to bf16
// before PV, output bf16.
//
// One CTA per (batch row, KV head, tile of QT queries), QT * group_size = 128 rows: warp w owns 16 rows (one
// q head x 16 queries), Q fragments in registers, K/V tiles of 64 tokens staged through shared memory with
// slot-table indirect 16-byte loads and shared by all the group's heads (GQA reuse), mma.sync m16n8k16
// bf16 / fp32 for S = Q K^T and O += P V (P goes from the S accumulators straight into A fragments).
// This is the step on the other side of the decode path; it is a correct tensor-core kernel,
// not yet a tuned one (no TMA pipeline: prefill is outside the measured hot path).
#include "common.cuh"
#include "launch.h"

namespace svb200 {

constexpr int kPaThreads = 256;
constexpr int kPaKTile = 64;
constexpr int kPaStride = kRowBytes + 16;
constexpr float kPaLog2e = 1.4426950408889634f;

struct PrefillAttnArgs {
  const __nv_bfloat16* q;  // [T, Hq, 128]
  int64_t q_stride;
  const __nv_bfloat16* k_cache;  // [slots, Hkv, 128]
  const __nv_bfloat16* v_cache;
  int64_t num_slots;
  const int32_t* slot_table;
  int64_t table_stride;
  const int32_t* req_rows;
  const int32_t* start_loc;          // [B] first chunk token of the row inside q / out
  const int32_t* context_lens;       // [B] physical row length including the chunk
  const int32_t* prompt_cache_lens;  // [B]
  __nv_bfloat16* out;                // [T, Hq, 128]
  int64_t out_stride;
  int hq, hkv, qt;  // qt: queries per CTA tile (a multiple of 16)
  float scale_log2;
};

__global__ void __launch_bounds__(kPaThreads) prefill_attn_kernel(PrefillAttnArgs a) {
  __shared__ __align__(16) unsigned char k_tile[kPaKTile * kPaStride];
  __shared__ __align__(16) unsigned char v_tile[kPaKTile * kPaStride];
  pdl_wait();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, g = lane >> 2, t = lane & 3;
  const int gsz = a.hq / a.hkv;
  const int b = blockIdx.y / a.hkv, kvh = blockIdx.y % a.hkv;
  const int ctx = a.context_lens[b], cache_len = a.prompt_cache_lens[b];
  const int chunk_len = ctx - cache_len;
  const int q0 = blockIdx.x * a.qt;  // first query (chunk-relative) of this tile
  if (q0 >= chunk_len) return;
  const int32_t* table = a.slot_table + (int64_t)a.req_rows[b] * a.table_stride;
  const int64_t tok_base = a.start_loc[b];

  // ---- this warp's 16 rows: head hl of the group, queries q0 + qi*16 .. +15
  const int tiles_q = a.qt / 16;
  const int hl = warp / tiles_q, qi = warp - hl * tiles_q;
  const bool warp_ok = hl < gsz;
  const int head = kvh * gsz + min(hl, gsz - 1);
  int row_rel[2];
  uint32_t qa[8][4];
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    const int r = q0 + qi * 16 + g + half * 8;
    row_rel[half] = (warp_ok && r < chunk_len) ? r : -1;
    const uint32_t* qrow = reinterpret_cast<const uint32_t*>(a.q + (tok_base + min(r, chunk_len - 1)) * a.q_stride +
                                                             (int64_t)head * kHeadDim);
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
      qa[kk][half] = row_rel[half] >= 0 ? qrow[kk * 8 + t] : 0u;
      qa[kk][2 + half] = row_rel[half] >= 0 ? qrow[kk * 8 + 4 + t] : 0u;
    }
  }
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  float o_acc[16][4];
#pragma unroll
  for (int nt = 0; nt < 16; ++nt) o_acc[nt][0] = o_acc[nt][1] = o_acc[nt][2] = o_acc[nt][3] = 0.f;

  // keys any query of this tile may see: positions < min(q0 + qt, chunk_len) + cache_len
  const int kv_end = min(q0 + a.qt, chunk_len) + cache_len;
  for (int kv0 = 0; kv0 < kv_end; kv0 += kPaKTile) {
    __syncthreads();
    for (int c = tid; c < kPaKTile * 16; c += kPaThreads) {
      const int tok = c >> 4, seg = c & 15;
      const int kv = kv0 + tok;
      uint4 kx = make_uint4(0, 0, 0, 0), vx = kx;
      if (kv < kv_end) {
        const int slot = table[kv];
        if (slot >= 0 && slot < a.num_slots) {
          const int64_t off = ((int64_t)slot * a.hkv + kvh) * kHeadDim + seg * 8;
          kx = *reinterpret_cast<const uint4*>(a.k_cache + off);
          vx = *reinterpret_cast<const uint4*>(a.v_cache + off);
        }
      }
      *reinterpret_cast<uint4*>(k_tile + tok * kPaStride + seg * 16) = kx;
      *reinterpret_cast<uint4*>(v_tile + tok * kPaStride + seg * 16) = vx;
    }
    __syncthreads();
    if (!warp_ok) continue;
    // the warp's rows see nothing beyond their own position: skip sub-tiles wholly in the future
    const int warp_last = min(q0 + qi * 16 + 15, chunk_len - 1) + cache_len;
    for (int sub = 0; sub < kPaKTile / 16 && kv0 + sub * 16 <= warp_last; ++sub) {
      float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      const uint32_t kaddr = smem_u32(k_tile) + (sub * 16 + (lane & 7) + (lane >> 4) * 8) * kPaStride + ((lane >> 3) & 1) * 16;
#pragma unroll
      for (int kk = 0; kk < 8; ++kk) {
        uint32_t kb[4];
        ldmatrix_x4(kb, kaddr + kk * 32);
        mma_16816(sc[0], qa[kk], kb[0], kb[1]);
        mma_16816(sc[1], qa[kk], kb[2], kb[3]);
      }
      // mask (causal, row validity), scale into log2 units
      float p[2][4];
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const int kv = kv0 + sub * 16 + nt * 8 + 2 * t + (e & 1);
          const int r = row_rel[e >> 1];
          p[nt][e] = (r >= 0 && kv <= r + cache_len && kv < kv_end) ? sc[nt][e] * a.scale_log2 : -INFINITY;
        }
      float alpha[2];
#pragma unroll
      for (int half = 0; half < 2; ++half) {
        float mx = fmaxf(fmaxf(p[0][2 * half], p[0][2 * half + 1]), fmaxf(p[1][2 * half], p[1][2 * half + 1]));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        const float mnew = fmaxf(m_run[half], mx);
        const float msafe = mnew == -INFINITY ? 0.f : mnew;
        alpha[half] = fast_exp2(m_run[half] - msafe);
        float sum = 0.f;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const float v = fast_exp2(p[nt][2 * half + e] - msafe);
            p[nt][2 * half + e] = v;
            sum += v;
          }
        l_run[half] = l_run[half] * alpha[half] + sum;  // per-lane partial sums: reduced over t at the end
        m_run[half] = mnew;
      }
#pragma unroll
      for (int nt = 0; nt < 16; ++nt) {
        o_acc[nt][0] *= alpha[0];
        o_acc[nt][1] *= alpha[0];
        o_acc[nt][2] *= alpha[1];
        o_acc[nt][3] *= alpha[1];
      }
      // P (16 rows x 16 tokens) in bf16: the S accumulators are already laid out like an A fragment
      uint32_t pa[4];
      pa[0] = pack_bf16x2(p[0][0], p[0][1]);
      pa[1] = pack_bf16x2(p[0][2], p[0][3]);
      pa[2] = pack_bf16x2(p[1][0], p[1][1]);
      pa[3] = pack_bf16x2(p[1][2], p[1][3]);
      // V^T fragments: B[k = token][n = dim] through ldmatrix.trans on V rows
      const uint32_t vaddr = smem_u32(v_tile) + (sub * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * kPaStride + (lane >> 4) * 16;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) {
        uint32_t vb[4];  // {b0, b1} of dims dt*16 .. +7, {b0, b1} of dims dt*16+8 .. +15
        ldmatrix_x4_trans(vb, vaddr + dt * 32);
        mma_16816(o_acc[2 * dt], pa, vb[0], vb[1]);
        mma_16816(o_acc[2 * dt + 1], pa, vb[2], vb[3]);
      }
    }
  }
  if (!warp_ok) return;
#pragma unroll
  for (int half = 0; half < 2; ++half) {
    float l = l_run[half];
    l += __shfl_xor_sync(0xffffffffu, l, 1);
    l += __shfl_xor_sync(0xffffffffu, l, 2);
    if (row_rel[half] < 0) continue;
    const float inv = l > 0.f ? 1.f / l : 0.f;
    uint32_t* orow = reinterpret_cast<uint32_t*>(a.out + (tok_base + row_rel[half]) * a.out_stride + (int64_t)head * kHeadDim);
#pragma unroll
    for (int nt = 0; nt < 16; ++nt)
      orow[nt * 4 + t] = pack_bf16x2(o_acc[nt][2 * half] * inv, o_acc[nt][2 * half + 1] * inv);
  }
}

int launch_prefill_attn(const PrefillAttnHostArgs& h, cudaStream_t stream) {
  if (h.batch <= 0 || h.max_chunk_len <= 0) return 0;
  const int gsz = h.hq / h.hkv;
  int qt = (kPaThreads / 32) / gsz * 16;  // 8 warps: qt/16 query tiles x gsz heads
  if (qt < 16) qt = 16;
  PrefillAttnArgs a{};
  a.q = (const __nv_bfloat16*)h.q; a.q_stride = h.q_stride; a.k_cache = (const __nv_bfloat16*)h.k_cache;
  a.v_cache = (const __nv_bfloat16*)h.v_cache; a.num_slots = h.num_slots; a.slot_table = h.slot_table;
  a.table_stride = h.table_stride; a.req_rows = h.req_rows; a.start_loc = h.start_loc; a.context_lens = h.context_lens;
  a.prompt_cache_lens = h.prompt_cache_lens; a.out = (__nv_bfloat16*)h.out; a.out_stride = h.out_stride;
  a.hq = h.hq; a.hkv = h.hkv; a.qt = qt; a.scale_log2 = h.sm_scale * kPaLog2e;
  dim3 grid((h.max_chunk_len + qt - 1) / qt, h.batch * h.hkv);
  const cudaError_t e = launch_kernel(prefill_attn_kernel, grid, dim3(kPaThreads), 0, stream, a);
  if (e != cudaSuccess) return cuda_error("prefill_attn launch", e);
  return check_launch("prefill_attn");
}

}  // namespace svb200
