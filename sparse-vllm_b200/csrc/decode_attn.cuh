// Sparse paged GQA decode attention for sm_100a: device code (the stand-alone kernel is in decode_attn.cu,
// the launch fused with the Quest selection in quest.cu).
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/kernels/triton/):
//   gqa_flash_decoding_stage1.py:6-121 (stage 1), 211-295 (2D head-max score),
//   flash_decoding_stage2.py:8-46 (combine).  Conventions kept: logits scaled by
//   head_dim^-0.5, softmax numerators cast to bf16 before PV, per-partial
//   normalisation by the partial's own sum, lse = max + log(sum), empty row -> 0.
//
// Decomposition (stream-K over token pages instead of the reference's fixed
// BLOCK_SEQ grid): a "row" is (batch entry b, head group hg of HPC KV heads); a
// "unit" is one 16-token page of a row.  All units of the launch form one
// sequence that is cut into gridDim.x contiguous ranges, one per persistent CTA,
// whatever the batch and the per-row lengths are.  The ranges are equal in COST,
// not in units: a CTA whose range holds a row boundary pays a second epilogue, a
// q reload, a pipeline bubble and -- arriving last for that reason -- the row's
// merge (~10 us at cfg5, where they were the launch's tail), so the cut is made
// in a virtual sequence with `split_pen` padding units at every row boundary:
// such a CTA streams split_pen units less, and a cut point that falls into the
// padding becomes a cut exactly at the row boundary (no split at all).  A CTA
// whose range covers a whole row writes bf16 output directly; otherwise it
// publishes one fp32 partial (at most two per CTA) and whoever completes the
// row's arrival ticket merges them.
//
// Inside a CTA: warps 8-11 are producers.  A producer reads 16 token slots (prefetched kSlotPrefetch
// units ahead) and issues bulk async copies (cp.async.bulk, SASS UBLKCP) into a ring of S stages whose
// mbarriers count the bytes (S * stage bytes ~ 160-190 KB in flight per SM): lanes 0-15 copy K, lanes
// 16-31 V.  A copy moves a *granule* of GRAN consecutive token rows when their slots are consecutive
// (a Quest page is 16 consecutive slots) and the CTA stages every local KV head (HPC == Hkv: token rows
// are then contiguous in the cache), else one token row (HPC*256 bytes).  GRAN is chosen so that a copy
// is >= 1 KB: the copy engine issues one bulk copy per ~30-45 cycles per SM whatever its size, so 256-byte
// copies (one KV head per rank under TP 8) capped the launch at 1.6 TB/s.  Granules sit 16 bytes apart
// modulo 128 in shared memory and the mma tile's rows are a fixed permutation of the unit's tokens
// (row_tok) so that ldmatrix still hits distinct banks (two-way conflicts at GRAN 4).
// Warps 0-7 are consumers: warp (h, tl) owns KV head h and every TL-th unit: S^T(16 tok x 8 heads) =
// K Q^T on tensor cores (mma.sync m16n8k16 bf16, fp32 accumulate), fp32 online softmax with warp
// shuffles, O^T(128 x 8) += V^T P^T.
//
// Tensor parallelism (push epilogue): with a svb200_comm_t the final bf16 rows are also written into
// every rank's gathered [B][world*Hq][128] buffer over NVLink (peer stores), and the last CTA of the
// launch signals the peers and waits for theirs: when the launch completes, the all-gather of the head
// shards has completed as well (no NCCL launch per layer).
#pragma once

#include <type_traits>

#include "common.cuh"
#include "launch.h"

namespace svb200 {


constexpr int kConsumerWarps = 8;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kProducerWarps = 4;
constexpr int kSlotPrefetch = 6;
constexpr int kAttnThreads = kConsumerThreads + kProducerWarps * 32;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxSmem = 227 * 1024;

template <int HPC>
struct AttnGeom {
  static constexpr int TL = kConsumerWarps / HPC;    // consumer warps per KV head
  static constexpr int TOK_BYTES = HPC * kRowBytes;   // one token's head-group row
  static constexpr int GRAN = HPC == 1 ? 4 : 2;  // token rows per copy granule (1 KB at one or two heads, 2-4 KB above)
  static constexpr int GSTRIDE = GRAN * TOK_BYTES + 16;  // +16: granules (ldmatrix rows) hit distinct banks
  static constexpr int KV_BYTES = (kPage / GRAN) * GSTRIDE;  // one tensor of one stage
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;    // K then V
  // byte offset of token `tok` (0..15) inside a stage's K (or V) part
  __device__ static __forceinline__ int tok_off(int tok) { return (tok / GRAN) * GSTRIDE + (tok % GRAN) * TOK_BYTES; }
  // token held by row r of the 16-row mma tile: rows 0-7 (one ldmatrix phase) come from distinct granules
  __device__ static __forceinline__ int row_tok(int r) {
    return GRAN == 1 ? r : (GRAN == 2 ? 2 * (r & 7) + (r >> 3) : 4 * (r & 3) + (r >> 2));
  }
};

// consumer-only barrier (the producer warps never join)
template <int BAR>
__device__ __forceinline__ void consumer_sync() {
  asm volatile("bar.sync %0, %1;" ::"n"(BAR), "n"(kConsumerThreads) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// CTA that owns (virtual) unit u when U units are cut into G ranges [U*c/G, U*(c+1)/G).
__device__ __forceinline__ int unit_owner(int64_t u, int64_t U, int64_t G) {
  return (int)(((u + 1) * G + U - 1) / U - 1);
}
// The cut.  Row r's units sit at virtual positions [prefix[r] + pen*r, prefix[r+1] + pen*r), followed by pen
// padding positions (none after the last row).  to_unit: virtual position -> unit (padding maps to the next
// row's first unit).
struct AttnCut {
  const int* prefix;  // exclusive prefix of the rows' unit counts, R + 1 entries
  int R, pen;
  int64_t U, UV, G;   // units, virtual length, ranges
  bool small;         // UV * G fits 31 bits: 32-bit divisions (a 64-bit one is ~100 instructions on the start-up path)
  __device__ __forceinline__ int64_t vbegin(int r) const { return prefix[r] + (int64_t)pen * r; }
  __device__ __forceinline__ int64_t vstart(int c) const {  // first virtual position of CTA c
    return small ? (int64_t)((uint32_t)UV * (uint32_t)c / (uint32_t)G) : UV * c / G;
  }
  __device__ __forceinline__ int unit_at(int r, int64_t v) const {
    return prefix[r] + (int)min(v - vbegin(r), (int64_t)(prefix[r + 1] - prefix[r]));
  }
  __device__ __forceinline__ int row_of_virtual(int64_t v) const {  // largest r with vbegin(r) <= v
    int lo = 0, hi = R;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (vbegin(mid) <= v) lo = mid; else hi = mid;
    }
    return lo;
  }
  __device__ __forceinline__ int to_unit(int64_t v) const { return unit_at(row_of_virtual(v), v); }
  __device__ __forceinline__ int begin_of(int c) const { return to_unit(vstart(c)); }  // first unit of CTA c
  // [u0, u1) of CTA c and the row u0 lies in: the two searches run in one loop (two independent chains)
  __device__ __forceinline__ void range_of(int c, int& u0, int& u1, int& row0) const {
    const int64_t v0 = vstart(c), v1 = vstart(c + 1);
    int lo0 = 0, hi0 = R, lo1 = 0, hi1 = R;
    while (hi0 - lo0 > 1 || hi1 - lo1 > 1) {
      const int m0 = (lo0 + hi0) >> 1, m1 = (lo1 + hi1) >> 1;
      if (hi0 - lo0 > 1) { if (vbegin(m0) <= v0) lo0 = m0; else hi0 = m0; }
      if (hi1 - lo1 > 1) { if (vbegin(m1) <= v1) lo1 = m1; else hi1 = m1; }
    }
    u0 = unit_at(lo0, v0);
    u1 = c + 1 == G ? (int)U : unit_at(lo1, v1);
    row0 = lo0;
  }
  __device__ __forceinline__ int owner(int u, int r) const {  // CTA that owns unit u of row r
    const int64_t v = u + (int64_t)pen * r;
    if (small) return (int)((((uint32_t)v + 1u) * (uint32_t)G + (uint32_t)UV - 1u) / (uint32_t)UV) - 1;
    return unit_owner(v, UV, G);
  }
};

// ---- tensor-parallel push epilogue (see the file header)
// Called by ONE thread per CTA (every CTA of the grid, exactly once) after a CTA barrier that follows the CTA's
// last peer store.  System-scope fences are the expensive part of the hand-shake (a fence waits for the NVLink
// acknowledgements of everything outstanding), so there are exactly two on the critical path:
//   1. this thread's fence before its arrival: the barrier ordered the CTA's peer stores before it, the fence
//      makes them visible at system scope before the arrival (the writers themselves do not fence);
//   2. the last CTA's fence after the arrival counter told it that every CTA has arrived: it orders all of
//      their stores before the flags, which are then plain relaxed stores, one per peer (a release store per
//      peer would be a fence per peer).
// The last CTA then waits for the peers' flags: the launch does not complete before every rank's rows have
// landed here.  Flags only grow, peers can be at most one launch ahead.
static __device__ void push_finish(const PushArgs& ps) {
  int32_t* local = ps.sync[ps.rank];
  __threadfence_system();
  if (atomicAdd(local + kSyncDone, 1) != (int)gridDim.x - 1) return;
  local[kSyncDone] = 0;
  const int32_t e = local[kSyncEpoch] + 1;
  local[kSyncEpoch] = e;
  __threadfence_system();
  for (int r = 0; r < ps.world; ++r)
    if (r != ps.rank) st_relaxed_sys(ps.sync[r] + ps.rank, e);
  for (int r = 0; r < ps.world; ++r) {
    if (r == ps.rank) continue;
    SpinGuard guard;
    while (ld_acquire_sys(local + r) - e < 0) {
      __nanosleep(32);
      guard.tick(kPeerSpinLimitNs);
    }
  }
}

// ---- flag-in-data variant of the push (PushArgs::ll_words > 0; svb200_comm_t::peer_ll)
// Every 4 payload bytes (two bf16) travel as one 8-byte word {payload, epoch}; a 64-bit store is visible atomically,
// so nothing has to order data before a flag: the senders do not fence, there is no flag, and a receiver that reads
// this launch's epoch in a word has its payload.  Peers write into THIS rank's staging; every CTA of the launch,
// when its own attention work is done, polls its share of the staging and copies the payload into the local
// gathered tensor, so the launch still completes only when every rank's rows are here.  The epoch is the sync
// block's (bumped by the launch's last CTA, local atomics only); two staging areas alternate by its parity (a peer
// can be one launch ahead).  This takes the two system-scope fences and the flag flight off the critical path.
__device__ __forceinline__ int32_t push_epoch(const PushArgs& ps) {
  return *reinterpret_cast<const volatile int32_t*>(ps.sync[ps.rank] + kSyncEpoch) + 1;
}
__device__ __forceinline__ void st_ll2(unsigned long long* dst, uint32_t a, uint32_t b, int32_t e) {
  const unsigned long long hi = (unsigned long long)(uint32_t)e << 32;
  asm volatile("st.relaxed.sys.global.v2.u64 [%0], {%1, %2};" ::"l"(dst), "l"(hi | a), "l"(hi | b) : "memory");
}
// elements [d0, d0 + 8) of this rank's row (b, local head h) -> every rank's gathered tensor (d0 % 8 == 0)
__device__ __forceinline__ void push_row8(const PushArgs& ps, int batch, int hq, int32_t e, int b, int h, int d0, uint4 v) {
  const int64_t off = (int64_t)b * ps.stride_b + ((int64_t)ps.rank * hq + h) * kHeadDim + d0;
  if (ps.ll_words == 0) {
    for (int r = 0; r < ps.world; ++r) *reinterpret_cast<uint4*>(ps.out[r] + off) = v;
    return;
  }
  *reinterpret_cast<uint4*>(ps.out[ps.rank] + off) = v;
  const int64_t word = (((int64_t)ps.rank * batch + b) * hq + h) * (kHeadDim / 2) + (d0 >> 1) + (int64_t)(e & 1) * ps.ll_words;
  for (int r = 0; r < ps.world; ++r)
    if (r != ps.rank) {
      st_ll2(ps.ll[r] + word, v.x, v.y, e);
      st_ll2(ps.ll[r] + word + 2, v.z, v.w, e);
    }
}
// elements [d0, d0 + 4) (d0 % 4 == 0)
__device__ __forceinline__ void push_row4(const PushArgs& ps, int batch, int hq, int32_t e, int b, int h, int d0, uint2 v) {
  const int64_t off = (int64_t)b * ps.stride_b + ((int64_t)ps.rank * hq + h) * kHeadDim + d0;
  if (ps.ll_words == 0) {
    for (int r = 0; r < ps.world; ++r) *reinterpret_cast<uint2*>(ps.out[r] + off) = v;
    return;
  }
  *reinterpret_cast<uint2*>(ps.out[ps.rank] + off) = v;
  const int64_t word = (((int64_t)ps.rank * batch + b) * hq + h) * (kHeadDim / 2) + (d0 >> 1) + (int64_t)(e & 1) * ps.ll_words;
  for (int r = 0; r < ps.world; ++r)
    if (r != ps.rank) st_ll2(ps.ll[r] + word, v.x, v.y, e);
}
// End of a CTA's work, called by its 256 consumer threads (every CTA of the grid, exactly once; BAR: their barrier).
template <int BAR>
__device__ __forceinline__ void push_complete(const PushArgs& ps, int batch, int hq, int32_t e) {
  const int tid = threadIdx.x;
  if (ps.ll_words == 0) {
    if (tid == 0) push_finish(ps);  // (a consumer barrier ordered the CTA's peer stores before this)
    return;
  }
  const int64_t groups = ps.ll_words / 4;  // 4 words = 8 elements = one 16-byte piece of the gathered tensor
  const int64_t per_src = groups / ps.world;
  const unsigned long long* mine = ps.ll[ps.rank] + (int64_t)(e & 1) * ps.ll_words;
  for (int64_t g = (int64_t)blockIdx.x * kConsumerThreads + tid; g < groups; g += (int64_t)gridDim.x * kConsumerThreads) {
    const int src = (int)(g / per_src);
    if (src == ps.rank) continue;
    unsigned long long w0, w1, w2, w3;
    SpinGuard guard;
    while (true) {
      asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w0), "=l"(w1) : "l"(mine + 4 * g) : "memory");
      asm volatile("ld.relaxed.sys.global.v2.u64 {%0, %1}, [%2];" : "=l"(w2), "=l"(w3) : "l"(mine + 4 * g + 2) : "memory");
      if ((int32_t)(w0 >> 32) == e && (int32_t)(w1 >> 32) == e && (int32_t)(w2 >> 32) == e && (int32_t)(w3 >> 32) == e) break;
      __nanosleep(100);
      guard.tick(kPeerSpinLimitNs);
    }
    const int64_t rem = g - (int64_t)src * per_src;  // (b * hq + h) * 16 + piece
    const int row = (int)(rem >> 4), piece = (int)(rem & 15);
    const int b = row / hq, h = row - b * hq;
    *reinterpret_cast<uint4*>(ps.out[ps.rank] + (int64_t)b * ps.stride_b + ((int64_t)src * hq + h) * kHeadDim + piece * 8) =
        make_uint4((uint32_t)w0, (uint32_t)w1, (uint32_t)w2, (uint32_t)w3);
  }
  consumer_sync<BAR>();
  if (tid == 0) {
    int32_t* local = ps.sync[ps.rank];
    __threadfence();
    if (atomicAdd(local + kSyncDone, 1) == (int)gridDim.x - 1) {
      local[kSyncDone] = 0;
      local[kSyncEpoch] = e;
    }
  }
}

struct AttnSmemLayout {
  int ring_off, valid_off, bar_off, q_off, merge_off, prefix_off, entry_off, total;
};

template <int HPC, int NT>
__host__ __device__ inline AttnSmemLayout attn_smem_layout(int stages, int gsz, int rows, int batch) {
  using G = AttnGeom<HPC>;
  AttnSmemLayout L;
  int off = 0;
  L.ring_off = off;   off += stages * G::STAGE_BYTES;
  L.q_off = off;      off += HPC * gsz * kRowBytes;
  L.merge_off = off;  off += (G::TL > 1) ? (kConsumerWarps - HPC) * 32 * NT * 36 * 4 : 0;
  L.bar_off = off;    off += 2 * stages * 8;
  L.valid_off = off;  off += ((stages * 4 + 15) / 16) * 16;
  L.prefix_off = off; off += ((rows + 1) * 4 + 15) / 16 * 16;
  L.entry_off = off;  off += (2 * batch * 4 + 15) / 16 * 16;  // clamped length, table row of each batch entry
  L.total = off;
  return L;
}

// Set-up of a CTA's attention state, run by its 256 consumer threads (warps 0-7): ring barriers, the batch
// entries' clamped lengths and table rows, the exclusive prefix of the rows' unit counts, zero rows for empty
// entries.  len_of(b) = tokens of batch entry b this launch attends to.  The caller puts a barrier between this
// and attn_main (all 12 attention warps).
template <int HPC, int NT, class LenOf>
__device__ __forceinline__ void attn_setup(const AttnParams& p, unsigned char* smem, LenOf len_of) {
  using G = AttnGeom<HPC>;
  constexpr int TL = G::TL;
  const int gsz = p.hq / p.hkv;
  const int HG = p.hkv / HPC;
  const int R = p.batch * HG;
  const int S = p.stages;
  const AttnSmemLayout L = attn_smem_layout<HPC, NT>(S, gsz, R, p.batch);
  unsigned char* ring = smem + L.ring_off;
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem + L.q_off);
  float* sMerge = reinterpret_cast<float*>(smem + L.merge_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* empty_bar = full_bar + S;
  uint32_t* sValid = reinterpret_cast<uint32_t*>(smem + L.valid_off);
  int* sPrefix = reinterpret_cast<int*>(smem + L.prefix_off);
  int* sLen = reinterpret_cast<int*>(smem + L.entry_off);  // clamped length of batch entry b
  int* sTab = sLen + p.batch;                              // its row of the slot table
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool push = p.push.world > 1;
  (void)ring; (void)sQ; (void)sMerge; (void)sValid; (void)TL;
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);  // the producer's expect_tx arrival; the copies' bytes complete the phase
      mbar_init(&empty_bar[s], HPC);
    }
    mbar_fence_init();
  }
  pdl_wait();
  // the producers' first slot fetch then only depends on shared memory (one DRAM round trip less before the
  // first copy is issued)
  if (warp >= 1)
    for (int b = tid - 32; b < p.batch; b += kConsumerThreads - 32) {
      sLen[b] = len_of(b);
      sTab[b] = p.req_rows ? p.req_rows[b] : b;  // (a one-launch layer reads row b of the slots it packed)
    }
  if (warp == 0) {
    int carry = 0;
    for (int base = 0; base < R; base += 32) {
      const int r = base + lane;
      int n = 0;
      if (r < R) {
        const int len = len_of(r / HG);
        n = len > 0 ? (len + kPage - 1) / kPage : 0;
      }
      int inc = n;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += v;
      }
      if (r < R) sPrefix[r] = carry + inc - n;
      carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) sPrefix[R] = carry;
  }
  // rows without tokens produce zeros (reference: empty split -> 0 / -inf)
  for (int b = blockIdx.x; b < p.batch; b += gridDim.x) {
    if (len_of(b) <= 0) {
      uint4* o4 = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) +
                                           (int64_t)b * p.out_stride_b);
      for (int i = tid; i < p.hq * 16; i += kConsumerThreads) o4[i] = make_uint4(0, 0, 0, 0);
      if (push) {
        const int32_t e = push_epoch(p.push);
        for (int i = tid; i < p.hq * 16; i += kConsumerThreads)
          push_row8(p.push, p.batch, p.hq, e, b, i >> 4, (i & 15) * 8, make_uint4(0, 0, 0, 0));
      }
    }
  }
}

// The attention work of one CTA after attn_setup: warps 0-7 consume, warps 8-11 produce (other warps must not
// call it).  BAR: named barrier of the 256 consumer threads.  LOWREG: the caller runs under a 128-register cap
// (smaller batches of loads in the merge).  wait_row(b): called by a producer warp before it
// reads slots of batch entry b (a fused launch waits there for the entry's selection; stand-alone: nothing).
template <int HPC, int NT, bool WITH_SCORE, int BAR, bool LOWREG, class WaitRow>
__device__ __forceinline__ void attn_main(const AttnParams& p, unsigned char* smem, WaitRow wait_row,
                                          unsigned long long* dbg) {
  using G = AttnGeom<HPC>;
  constexpr int TL = G::TL;
  const int gsz = p.hq / p.hkv;
  const int HG = p.hkv / HPC;
  const int R = p.batch * HG;
  const int S = p.stages;
  const AttnSmemLayout L = attn_smem_layout<HPC, NT>(S, gsz, R, p.batch);
  unsigned char* ring = smem + L.ring_off;
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem + L.q_off);
  float* sMerge = reinterpret_cast<float*>(smem + L.merge_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* empty_bar = full_bar + S;
  uint32_t* sValid = reinterpret_cast<uint32_t*>(smem + L.valid_off);
  int* sPrefix = reinterpret_cast<int*>(smem + L.prefix_off);
  int* sLen = reinterpret_cast<int*>(smem + L.entry_off);  // clamped length of batch entry b
  int* sTab = sLen + p.batch;                              // its row of the slot table
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool push = p.push.world > 1;
  __shared__ int s_flag, s_defer;
  if (tid == 0) s_defer = 0;  // (read after several consumer barriers)
  const int64_t U = sPrefix[R];
  // never more ranges than units.  The contributors of a row are the CTAs owner(first unit) .. owner(last unit):
  // every CTA between them has units of the row (the padding is at most half a range long, so no range of
  // theirs is empty).
  const int64_t G0 = p.grid_cap > 0 && p.grid_cap < (int)gridDim.x ? p.grid_cap : (int)gridDim.x;
  const int64_t Gn = U < G0 ? U : G0;
  AttnCut cut;
  cut.prefix = sPrefix, cut.R = R, cut.U = U, cut.G = Gn;
  cut.pen = Gn > 0 ? (int)min((int64_t)p.split_pen, U / Gn / 2) : 0;
  cut.UV = U + (int64_t)cut.pen * (R - 1);
  cut.small = (cut.UV + 1) * (Gn + 1) < (int64_t)1 << 31;
  int u0 = 0, u1 = 0, r_first = 0;
  if ((int64_t)blockIdx.x < Gn) cut.range_of(blockIdx.x, u0, u1, r_first);
  const int32_t epoch = push ? push_epoch(p.push) : 0;
  if (u0 >= u1) {
    if (push && warp < kConsumerWarps) push_complete<BAR>(p.push, p.batch, p.hq, epoch);
    return;
  }
  // first row of the range: the non-empty row that holds u0 (u0 may sit in the padding before the next row)
  while (sPrefix[r_first + 1] <= u0) ++r_first;

  const __nv_bfloat16* kc = reinterpret_cast<const __nv_bfloat16*>(p.k_cache);
  const __nv_bfloat16* vc = reinterpret_cast<const __nv_bfloat16*>(p.v_cache);
  const int64_t slot_elems = (int64_t)p.hkv * kHeadDim;

  if (warp >= kConsumerWarps) {
    // ============================================================== producers
    // Producer warp pw feeds every NP-th unit.  Lane t (and t + 16) holds the slot of token t, fetched
    // kSlotPrefetch units of this warp ahead so the table's DRAM latency hides behind queued copies.
    // Unit i is fed by producer i % NP with NP | S: a stage is always refilled by the same warp, so
    // the empty-phase it waits for is never more than one ahead of the barrier (same parity
    // argument as for the consumers).
    const int n_units = u1 - u0;
    const int pw = warp - kConsumerWarps;
    const int NP = p.producers;
    if (pw >= NP) return;
    const int tok_lane = lane & 15;
    // prefetch cursor: row / batch entry of the unit whose slots are fetched next; the row's
    // length and table row are reloaded only when the cursor moves to another batch entry
    int r_pf = r_first, b_pf = r_first / HG, hg_pf = r_first % HG;
    wait_row(b_pf);
    int len_pf = sLen[b_pf];
    const int32_t* row_pf = p.slot_table + (int64_t)sTab[b_pf] * p.slot_table_stride;
    auto fetch = [&](int i) -> int {  // raw slot of token tok_lane in local unit i
      if (i >= n_units) return -1;
      const int u = u0 + i;
      bool moved = false;
      while (sPrefix[r_pf + 1] <= u) {
        ++r_pf;
        if (++hg_pf == HG) hg_pf = 0, ++b_pf, moved = true;
      }
      if (moved) {
        wait_row(b_pf);
        len_pf = sLen[b_pf];
        row_pf = p.slot_table + (int64_t)sTab[b_pf] * p.slot_table_stride;
      }
      const int t = (u - sPrefix[r_pf]) * kPage + tok_lane;
      if (t >= len_pf) return -1;
      return __ldg(row_pf + t);
    };
    const int64_t slot_bytes = slot_elems * 2;
    const uint64_t l2_policy = l2_evict_first_policy();
    // granule copies need the GRAN token rows contiguous in the cache: every local KV head staged together
    const bool gran_ok = G::GRAN > 1 && p.hkv == HPC && p.granule;
    int r_is = r_first, hg_is = r_first % HG;  // issue cursor
    int s = pw % S;
    uint32_t ph = (pw / S) & 1;  // phase parity of stage s at unit i
    // slots are fetched kSlotPrefetch units of this warp ahead: the table comes from DRAM and one
    // unit's copies are issued in far less than a DRAM round trip
    // pf is a circular FIFO indexed at compile time (the unit loop is unrolled kSlotPrefetch times) and a slot is
    // refilled only after its value was used: a register that waits for its table load is never copied, so the
    // load's latency is paid kSlotPrefetch units later and not at a register move right after the load.
    int pf[kSlotPrefetch];
#pragma unroll
    for (int d = 0; d < kSlotPrefetch; ++d) pf[d] = fetch(pw + d * NP);
    for (int i0 = pw; i0 < n_units; i0 += NP * kSlotPrefetch) {
#pragma unroll
     for (int d = 0; d < kSlotPrefetch; ++d) {
      const int i = i0 + d * NP;
      if (i >= n_units) break;
      const int slot_a = pf[d];
      const int my = (slot_a >= 0 && slot_a < p.num_slots) ? slot_a : -1;
      const uint32_t mask = __ballot_sync(0xffffffffu, my >= 0) & 0xffffu;
      while (sPrefix[r_is + 1] <= u0 + i) {
        ++r_is;
        if (++hg_is == HG) hg_is = 0;
      }
      // Bulk async copies (UBLKCP): lanes 0-15 copy K rows of tokens 0-15, lanes 16-31 the V rows, straight
      // from the lane that holds the slot (no shuffles of addresses); the bytes are counted on the stage's
      // mbarrier.  The first lane of a granule copies all GRAN rows when their slots are consecutive.
      // V rows of masked tokens are zeroed by hand (P is 0 there, but 0 x stale NaN would poison the
      // accumulator); K rows of masked tokens only feed logits that are masked.
      const bool is_v = lane >= 16;
      bool lead_bulk = false, covered = false;
      if constexpr (G::GRAN > 1) {
        bool contig = gran_ok && my >= 0;
#pragma unroll
        for (int x = 1; x < G::GRAN; ++x) {
          const int next = __shfl_down_sync(0xffffffffu, my, x);  // every lane takes part (no short-circuit around it)
          contig = contig && next == my + x;
        }
        const bool leader = (tok_lane % G::GRAN) == 0;
        const uint32_t gm = __ballot_sync(0xffffffffu, leader && contig);
        lead_bulk = leader && contig;
        covered = (gm >> (lane - tok_lane % G::GRAN)) & 1u;
      }
      const char* base = reinterpret_cast<const char*>(is_v ? vc : kc) + hg_is * G::TOK_BYTES;
      const uint32_t dst = smem_u32(ring + s * G::STAGE_BYTES) + (is_v ? G::KV_BYTES : 0) + G::tok_off(tok_lane);
      mbar_wait(&empty_bar[s], ph ^ 1);
      if (mask != 0xffffu) {
        if (my < 0 && is_v) {
          uint4* z = reinterpret_cast<uint4*>(ring + s * G::STAGE_BYTES + G::KV_BYTES + G::tok_off(tok_lane));
          for (int c = 0; c < G::TOK_BYTES / 16; ++c) z[c] = make_uint4(0, 0, 0, 0);
          fence_proxy_async();  // generic-proxy stores to a stage the async proxy (bulk copies) rewrites later
        }
        __syncwarp();  // the zeroed rows are ordered before lane 0's arrival below
      }
      if (lane == 0) {
        sValid[s] = mask;
        mbar_arrive_expect_tx(&full_bar[s], (uint32_t)__popc(mask) * 2u * (uint32_t)G::TOK_BYTES);
      }
      if (p.evict_first) {
        if (lead_bulk)
          bulk_g2s_hint(dst, base + (int64_t)my * slot_bytes, (uint32_t)(G::GRAN * G::TOK_BYTES), &full_bar[s], l2_policy);
        else if (my >= 0 && !covered)
          bulk_g2s_hint(dst, base + (int64_t)my * slot_bytes, (uint32_t)G::TOK_BYTES, &full_bar[s], l2_policy);
      } else {
        if (lead_bulk)
          bulk_g2s(dst, base + (int64_t)my * slot_bytes, (uint32_t)(G::GRAN * G::TOK_BYTES), &full_bar[s]);
        else if (my >= 0 && !covered)
          bulk_g2s(dst, base + (int64_t)my * slot_bytes, (uint32_t)G::TOK_BYTES, &full_bar[s]);
      }
      pf[d] = fetch(i + kSlotPrefetch * NP);
      s += NP;
      while (s >= S) s -= S, ph ^= 1;
     }
    }
    if (dbg && pw == 0 && lane == 0) dbg[4] = global_timer_ns();
    return;
  }

  // ================================================================ consumers
  const int g = lane >> 2, t = lane & 3;
  const int h_local = warp % HPC, tl = warp / HPC;
  const float scale_log2 = p.sm_scale * kLog2e;
  const int heads_row = HPC * gsz;

  // Merge of a row that several CTAs contributed to (run by all consumer warps of the CTA whose arrival
  // completed the row's ticket).
  auto merge_row = [&](int mr) {
    const int row_begin = sPrefix[mr], row_end = sPrefix[mr + 1];
    const int b = mr / HG, hg = mr - b * HG;
    const int c_first = cut.owner(row_begin, mr), c_last = cut.owner(row_end - 1, mr);
    const unsigned long long t_cmb = dbg ? global_timer_ns() : 0;
    // One warp per group of q heads, lane = 4 output dims.  Fast path (the usual case: a row has a handful of
    // contributors): lane c fetches the lse of contributor c AND every lane fetches its 4 dims of every
    // contributor's partial at once, so the merge costs ONE L2 round trip; the weights are applied when both
    // have arrived.  HB heads x NCMAX contributors x 4 floats are in flight per lane (96 registers: the
    // accumulators of the main loop are dead by now).
    __threadfence();
    const int nc = c_last - c_first + 1;
    auto fast = [&](auto hb_tag, auto nc_tag) {
      constexpr int HBF = decltype(hb_tag)::value, NCMAX = decltype(nc_tag)::value;
      int ps = 0;
      if (lane < nc) ps = 2 * (c_first + lane) + (cut.begin_of(c_first + lane) >= row_begin ? 0 : 1);
      for (int h0 = warp * HBF; h0 < heads_row; h0 += kConsumerWarps * HBF) {
        float lse[HBF];
        float4 v[NCMAX][HBF];
#pragma unroll
        for (int x = 0; x < HBF; ++x)
          lse[x] = (lane < nc && h0 + x < heads_row) ? __ldcg(p.ws_lse + (int64_t)ps * heads_row + h0 + x) : -INFINITY;
#pragma unroll
        for (int kc2 = 0; kc2 < NCMAX; ++kc2) {
          const int psk = __shfl_sync(0xffffffffu, ps, kc2);
          const float4* src = reinterpret_cast<const float4*>(p.ws_o + ((int64_t)psk * heads_row + h0) * kHeadDim) + lane;
#pragma unroll
          for (int x = 0; x < HBF; ++x)
            v[kc2][x] = (kc2 < nc && h0 + x < heads_row) ? __ldcg(src + x * (kHeadDim / 4)) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int x = 0; x < HBF; ++x) {
          float cm = lse[x];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
          const float wgt = lse[x] == -INFINITY ? 0.f : __expf(lse[x] - cm);
          float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
          float wsum = 0.f;
#pragma unroll
          for (int kc2 = 0; kc2 < NCMAX; ++kc2) {
            const float wk = __shfl_sync(0xffffffffu, wgt, kc2);
            if (kc2 < nc && wk != 0.f) {
              acc.x += wk * v[kc2][x].x, acc.y += wk * v[kc2][x].y, acc.z += wk * v[kc2][x].z, acc.w += wk * v[kc2][x].w;
              wsum += wk;
            }
          }
          if (h0 + x < heads_row) {
            const float inv = wsum > 0.f ? 1.f / wsum : 0.f;
            uint2 o2;
            o2.x = pack_bf16x2(acc.x * inv, acc.y * inv);
            o2.y = pack_bf16x2(acc.z * inv, acc.w * inv);
            __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) + (int64_t)b * p.out_stride_b +
                                  (int64_t)(hg * heads_row + h0 + x) * kHeadDim;
            *reinterpret_cast<uint2*>(orow + lane * 4) = o2;
            if (push) push_row4(p.push, p.batch, p.hq, epoch, b, hg * heads_row + h0 + x, lane * 4, o2);
          }
        }
      }
    };
    bool merged = true;
    if (LOWREG && heads_row > 16 && nc <= 6) fast(std::integral_constant<int, 2>(), std::integral_constant<int, 6>());
    else if (heads_row > 16 && nc <= 6) fast(std::integral_constant<int, 4>(), std::integral_constant<int, 6>());
    else if (LOWREG && heads_row > 8 && heads_row <= 16 && nc <= 6) fast(std::integral_constant<int, 2>(), std::integral_constant<int, 6>());
    else if (LOWREG && heads_row <= 8 && nc <= 12) fast(std::integral_constant<int, 1>(), std::integral_constant<int, 12>());
    else if (!LOWREG && heads_row > 8 && heads_row <= 16 && nc <= 12) fast(std::integral_constant<int, 2>(), std::integral_constant<int, 12>());
    else if (!LOWREG && heads_row <= 8 && nc <= 24) fast(std::integral_constant<int, 1>(), std::integral_constant<int, 24>());
    else merged = false;
    // General path (any number of contributors): the lse first, then the partials that carry weight.
    constexpr int HB = 4;  // heads a warp merges at once: their loads are all independent
    for (int h0 = merged ? heads_row : warp * HB; h0 < heads_row; h0 += kConsumerWarps * HB) {
      float m[HB], wsum[HB];
      float4 acc[HB];
#pragma unroll
      for (int x = 0; x < HB; ++x) m[x] = -INFINITY, wsum[x] = 0.f, acc[x] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int c0 = 0; c0 < nc; c0 += 32) {
        const int c = c_first + c0 + lane;
        int ps = 0;
        float lse[HB], wgt[HB];
        if (c <= c_last) ps = 2 * c + (cut.begin_of(c) >= row_begin ? 0 : 1);
#pragma unroll
        for (int x = 0; x < HB; ++x)
          lse[x] = (c <= c_last && h0 + x < heads_row)
                       ? __ldcg(p.ws_lse + (int64_t)ps * heads_row + h0 + x)
                       : -INFINITY;
#pragma unroll
        for (int x = 0; x < HB; ++x) {
          float cm = lse[x];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) cm = fmaxf(cm, __shfl_xor_sync(0xffffffffu, cm, o));
          const float mnew = fmaxf(m[x], cm);
          const float rescale = m[x] == -INFINITY ? 0.f : __expf(m[x] - mnew);
          acc[x].x *= rescale, acc[x].y *= rescale, acc[x].z *= rescale, acc[x].w *= rescale;
          wsum[x] *= rescale;
          wgt[x] = lse[x] == -INFINITY ? 0.f : __expf(lse[x] - mnew);
          m[x] = mnew;
        }
        const int cnt = min(32, nc - c0);
#pragma unroll 4
        for (int kc2 = 0; kc2 < cnt; ++kc2) {
          const int psk = __shfl_sync(0xffffffffu, ps, kc2);
          const float4* src = reinterpret_cast<const float4*>(
                                  p.ws_o + ((int64_t)psk * heads_row + h0) * kHeadDim) + lane;
#pragma unroll
          for (int x = 0; x < HB; ++x) {
            const float wk = __shfl_sync(0xffffffffu, wgt[x], kc2);
            if (wk != 0.f) {
              const float4 v = __ldcg(src + x * (kHeadDim / 4));
              acc[x].x += wk * v.x, acc[x].y += wk * v.y, acc[x].z += wk * v.z, acc[x].w += wk * v.w;
              wsum[x] += wk;
            }
          }
        }
      }
#pragma unroll
      for (int x = 0; x < HB; ++x) {
        if (h0 + x < heads_row) {
          const float inv = wsum[x] > 0.f ? 1.f / wsum[x] : 0.f;
          uint2 o2;
          o2.x = pack_bf16x2(acc[x].x * inv, acc[x].y * inv);
          o2.y = pack_bf16x2(acc[x].z * inv, acc[x].w * inv);
          __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                (int64_t)b * p.out_stride_b +
                                (int64_t)(hg * heads_row + h0 + x) * kHeadDim;
          *reinterpret_cast<uint2*>(orow + lane * 4) = o2;
          if (push) push_row4(p.push, p.batch, p.hq, epoch, b, hg * heads_row + h0 + x, lane * 4, o2);
        }
      }
    }
    if (dbg && tid == 0) dbg[5] += global_timer_ns() - t_cmb;
  };

  int defer_row = -1, defer_total = 0, defer_old = -1;
  int r = r_first;
  int i = 0;  // index of the next unit inside this CTA's range
  while (u0 + i < u1) {
    while (sPrefix[r + 1] <= u0 + i) ++r;
    const int row_begin = sPrefix[r], row_end = sPrefix[r + 1];
    const int seg_begin = u0 + i, seg_end = min(u1, row_end);
    const int n = seg_end - seg_begin;
    const bool complete = seg_begin == row_begin && seg_end == row_end;
    const int b = r / HG, hg = r - b * HG;

    // q rows of this head group -> smem -> B fragments (k = head dim, n = q head)
    uint32_t qb[NT][8][2];
    {
      const uint4* qsrc = reinterpret_cast<const uint4*>(
          reinterpret_cast<const __nv_bfloat16*>(p.q) + (int64_t)b * p.q_stride_b +
          (int64_t)hg * heads_row * kHeadDim);
      uint4* qdst = reinterpret_cast<uint4*>(sQ);
      for (int c = tid; c < heads_row * 16; c += kConsumerThreads) qdst[c] = qsrc[c];
      consumer_sync<BAR>();
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
        for (int e = 0; e < 4; ++e) o_acc[nt][dt][e] = 0.f;
    }

    for (int j = i + ((tl - i % TL + TL) % TL); j < i + n; j += TL) {
      const int s = j % S;
      mbar_wait(&full_bar[s], (uint32_t)(j / S) & 1u);
      if (dbg && tid == 0 && j == 0) dbg[2] = global_timer_ns();
      const uint32_t mask = sValid[s];
      const bool tv0 = (mask >> G::row_tok(g)) & 1u, tv1 = (mask >> G::row_tok(g + 8)) & 1u;
      if (mask != 0u && !p.no_math) {
        const uint32_t kbase = smem_u32(ring + s * G::STAGE_BYTES) + h_local * kRowBytes;
        const uint32_t vbase = kbase + G::KV_BYTES;
        float sc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
        {
          const uint32_t kaddr =
              kbase + G::tok_off(G::row_tok((lane & 7) + ((lane >> 3) & 1) * 8)) + (lane >> 4) * 16;
#pragma unroll
          for (int kk = 0; kk < 8; ++kk) {
            uint32_t a[4];
            ldmatrix_x4(a, kaddr + kk * 32);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816(sc[nt], a, qb[nt][kk][0], qb[nt][kk][1]);
          }
        }
        if constexpr (WITH_SCORE) {
          // raw (unscaled) max over this KV head's q heads; KV heads meet in fp32 atomics
          float r0 = -INFINITY, r1 = -INFINITY;
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const int col = nt * 8 + 2 * t;
            if (col < gsz) r0 = fmaxf(r0, sc[nt][0]), r1 = fmaxf(r1, sc[nt][2]);
            if (col + 1 < gsz) r0 = fmaxf(r0, sc[nt][1]), r1 = fmaxf(r1, sc[nt][3]);
          }
          r0 = fmaxf(r0, __shfl_xor_sync(0xffffffffu, r0, 1));
          r0 = fmaxf(r0, __shfl_xor_sync(0xffffffffu, r0, 2));
          r1 = fmaxf(r1, __shfl_xor_sync(0xffffffffu, r1, 1));
          r1 = fmaxf(r1, __shfl_xor_sync(0xffffffffu, r1, 2));
          if (t == 0) {
            const int t0 = (seg_begin + (j - i) - row_begin) * kPage;
            float* srow = p.attn_score + (int64_t)b * p.score_stride + t0;
            if (tv0) atomic_max_f32(srow + G::row_tok(g), r0);
            if (tv1) atomic_max_f32(srow + G::row_tok(g + 8), r1);
          }
        }
        uint32_t pb[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float x[4];
          x[0] = tv0 ? sc[nt][0] * scale_log2 : -INFINITY;
          x[1] = tv0 ? sc[nt][1] * scale_log2 : -INFINITY;
          x[2] = tv1 ? sc[nt][2] * scale_log2 : -INFINITY;
          x[3] = tv1 ? sc[nt][3] * scale_log2 : -INFINITY;
          float alpha[2];
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            float mloc = fmaxf(x[e], x[e + 2]);
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 4));
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 8));
            mloc = fmaxf(mloc, __shfl_xor_sync(0xffffffffu, mloc, 16));
            const float mnew = fmaxf(m_run[nt][e], mloc);
            const float msafe = mnew == -INFINITY ? 0.f : mnew;
            alpha[e] = fast_exp2(m_run[nt][e] - msafe);
            x[e] = fast_exp2(x[e] - msafe);
            x[e + 2] = fast_exp2(x[e + 2] - msafe);
            l_run[nt][e] = l_run[nt][e] * alpha[e] + x[e] + x[e + 2];
            m_run[nt][e] = mnew;
          }
#pragma unroll
          for (int dt = 0; dt < 8; ++dt) {
            o_acc[nt][dt][0] *= alpha[0];
            o_acc[nt][dt][1] *= alpha[1];
            o_acc[nt][dt][2] *= alpha[0];
            o_acc[nt][dt][3] *= alpha[1];
          }
          // P (tok x head) -> bf16 -> transpose -> B fragment (k = token, n = head)
          pb[nt][0] = movmatrix_trans(pack_bf16x2(x[0], x[1]));
          pb[nt][1] = movmatrix_trans(pack_bf16x2(x[2], x[3]));
        }
        {
          const uint32_t vaddr =
              vbase + G::tok_off(G::row_tok((lane & 7) + (lane >> 4) * 8)) + ((lane >> 3) & 1) * 16;
#pragma unroll
          for (int dt = 0; dt < 8; ++dt) {
            uint32_t a[4];
            ldmatrix_x4_trans(a, vaddr + dt * 32);
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) mma_16816(o_acc[nt][dt], a, pb[nt][0], pb[nt][1]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }

    // ---------------------------------------------------- segment epilogue
    const unsigned long long t_epi = dbg ? global_timer_ns() : 0;
    // per-head sums live as per-lane partials: reduce over the 8 token lanes
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 2; ++e) {
        float l = l_run[nt][e];
        l += __shfl_xor_sync(0xffffffffu, l, 4);
        l += __shfl_xor_sync(0xffffffffu, l, 8);
        l += __shfl_xor_sync(0xffffffffu, l, 16);
        l_run[nt][e] = l;
      }
    if constexpr (TL > 1) {
      // warps tl > 0 hand their state to the tl == 0 warp of the same KV head
      if (tl > 0) {
        const int widx = (tl - 1) * HPC + h_local;
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) {
          float* rec = sMerge + ((widx * NT + nt) * 36) * 32 + lane;
          rec[0 * 32] = m_run[nt][0];
          rec[1 * 32] = m_run[nt][1];
          rec[2 * 32] = l_run[nt][0];
          rec[3 * 32] = l_run[nt][1];
#pragma unroll
          for (int dt = 0; dt < 8; ++dt)
#pragma unroll
            for (int e = 0; e < 4; ++e) rec[(4 + dt * 4 + e) * 32] = o_acc[nt][dt][e];
        }
      }
      consumer_sync<BAR>();
      if (tl == 0) {
        for (int o = 1; o < TL; ++o) {
          const int widx = (o - 1) * HPC + h_local;
#pragma unroll
          for (int nt = 0; nt < NT; ++nt) {
            const float* rec = sMerge + ((widx * NT + nt) * 36) * 32 + lane;
            float sa[2], sb[2];
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const float mo = rec[e * 32], lo = rec[(2 + e) * 32];
              const float mnew = fmaxf(m_run[nt][e], mo);
              const float msafe = mnew == -INFINITY ? 0.f : mnew;
              sa[e] = fast_exp2(m_run[nt][e] - msafe);
              sb[e] = fast_exp2(mo - msafe);
              l_run[nt][e] = l_run[nt][e] * sa[e] + lo * sb[e];
              m_run[nt][e] = mnew;
            }
#pragma unroll
            for (int dt = 0; dt < 8; ++dt)
#pragma unroll
              for (int e = 0; e < 4; ++e)
                o_acc[nt][dt][e] =
                    o_acc[nt][dt][e] * sa[e & 1] + rec[(4 + dt * 4 + e) * 32] * sb[e & 1];
          }
        }
      }
    }
    // publish: bf16 output when this CTA saw the whole row, else an fp32 partial
    const int pslot = 2 * blockIdx.x + (seg_begin == u0 ? 0 : 1);
    // a whole row: the bf16 rows are staged in sQ (every warp has its q fragments in registers by now: the
    // barrier makes sure of it) and leave as 16-byte vectors, to this rank and, under TP, to every peer
    if (complete) consumer_sync<BAR>();
    if (tl == 0) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt)
#pragma unroll
        for (int e = 0; e < 2; ++e) {
          const int hi = nt * 8 + 2 * t + e;
          if (hi < gsz) {
            const int head_local = h_local * gsz + hi;
            const float l = l_run[nt][e];
            const float inv = l > 0.f ? 1.f / l : 0.f;
            if (complete) {
              __nv_bfloat16* orow = sQ + head_local * kHeadDim;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = __float2bfloat16_rn(o_acc[nt][dt][e] * inv);
                orow[dt * 16 + g + 8] = __float2bfloat16_rn(o_acc[nt][dt][2 + e] * inv);
              }
            } else {
              float* orow = p.ws_o + ((int64_t)pslot * heads_row + head_local) * kHeadDim;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = o_acc[nt][dt][e] * inv;
                orow[dt * 16 + g + 8] = o_acc[nt][dt][2 + e] * inv;
              }
              if (g == 0)
                p.ws_lse[(int64_t)pslot * heads_row + head_local] =
                    l > 0.f ? m_run[nt][e] * kLn2 + __logf(l) : -INFINITY;
            }
          }
        }
    }
    if (complete) {
      consumer_sync<BAR>();
      const uint4* src = reinterpret_cast<const uint4*>(sQ);
      const int64_t head0 = (int64_t)hg * heads_row * kHeadDim;
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) + (int64_t)b * p.out_stride_b + head0);
      for (int c = tid; c < heads_row * 16; c += kConsumerThreads) dst[c] = src[c];
      if (push)
        for (int c = tid; c < heads_row * 16; c += kConsumerThreads)
          push_row8(p.push, p.batch, p.hq, epoch, b, hg * heads_row + (c >> 4), (c & 15) * 8, src[c]);
    }
    if (!complete) {
      // Arrival ticket of the row, counted in KV-head warps (HPC per contributing CTA): whoever completes it merges
      // every partial.  A segment that is not the CTA's last one (the tail of a row that began in another CTA)
      // arrives early and is practically never the last: its warps fire their arrival and go on with the next
      // row at once; the returned count is only looked at when the CTA is done (defer_*).  The CTA's last segment
      // arrives as a whole and merges right away if it completes the row.
      const int c_first = cut.owner(row_begin, r), c_last = cut.owner(row_end - 1, r);
      const int arrivals = (c_last - c_first + 1) * HPC;
      if (seg_end != u1) {
        if (tl == 0) {
          __threadfence();
          __syncwarp();
          if (lane == 0) defer_old = atomicAdd(&p.tickets[r], 1);
        }
        defer_row = r;
        defer_total = arrivals;
      } else {
        __threadfence();
        consumer_sync<BAR>();
        if (tid == 0) {
          const int ticket = atomicAdd(&p.tickets[r], HPC);
          s_flag = (ticket + HPC == arrivals);
          if (s_flag) p.tickets[r] = 0;  // scratch stays zeroed for the next launch / replay
        }
        consumer_sync<BAR>();
        if (s_flag) merge_row(r);
      }
    }
    consumer_sync<BAR>();  // sQ / sMerge / s_flag are reused by the next segment
    if (dbg && tid == 0) dbg[6] += global_timer_ns() - t_epi;
    i += n;
  }
  if (defer_row >= 0) {
    if (tl == 0 && lane == 0 && defer_old == defer_total - 1) {
      s_defer = 1;
      p.tickets[defer_row] = 0;
    }
    consumer_sync<BAR>();
    if (s_defer) {
      merge_row(defer_row);
      consumer_sync<BAR>();
    }
  }
  if (dbg && tid == 0) dbg[3] = global_timer_ns();
  if (push) push_complete<BAR>(p.push, p.batch, p.hq, epoch);  // (the loop's last consumer barrier ordered the CTA's peer stores before this)
}

}  // namespace svb200
