// Sparse paged GQA decode attention for sm_100a.
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/kernels/triton/):
//   gqa_flash_decoding_stage1.py:6-121 (stage 1), 211-295 (2D head-max score),
//   flash_decoding_stage2.py:8-46 (combine).  Conventions kept: logits scaled by
//   head_dim^-0.5, softmax numerators cast to bf16 before PV, per-partial
//   normalisation by the partial's own sum, lse = max + log(sum), empty row -> 0.
//
// Decomposition: a "row" is (batch entry b, head group hg of HPC KV heads); a "unit" is
// one 16-token page of a row.  Units are handed out dynamically (sched.cuh): gridDim/rows
// CTAs start on each row and claim guided batches of 2-unit chunks from the row's counter,
// then steal from the row with the most chunks left, so SMs the memory system serves faster
// attend more tokens (the reference cuts every row into fixed 256-token blocks).  A CTA keeps
// ONE running softmax state per visit of a row, whatever chunks it got (softmax accumulation
// is order-free).  A CTA that attended a whole row writes bf16 output directly; otherwise it
// publishes one normalised fp32 partial + lse into its reserved contributor slot of the row
// and adds its units to the row's counter; the CTA that completes the row merges the partials.
//
// Inside a CTA: a scheduler warp claims chunks; producer warps resolve the 32 token slots of a
// chunk (fetched kIdDepth chunks ahead) and copy each token's HPC*256 contiguous bytes of K and
// of V with 16-byte LDGSTS (a whole warp per 512 B, masked tokens zero-filled) into a ring of S
// stages whose mbarriers complete when the copies land (~200 KB in flight per SM).  (A first
// version issued one UBLKCP bulk copy per token row: the per-lane operands serialise through
// uniform registers at ~110 cycles per copy and capped the kernel at 2.5 TB/s, profiles/r1_b_*.)
// Warps 0-7 are consumers: warp (h, tl) owns KV head h and every TL-th unit:
// S^T(16 tok x 8 heads) = K Q^T on tensor cores, fp32 online softmax with warp shuffles,
// O^T(128 x 8) += V^T P^T.
#include "common.cuh"
#include "launch.h"
#include "sched.cuh"

namespace svb200 {

constexpr int kConsumerWarps = 8;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kProducerWarps = 3;  // + 1 scheduler warp
constexpr int kAttnIdDepth = 2;    // chunks whose token slots are in flight ahead of the copies
constexpr int kAttnThreads = kConsumerThreads + (kProducerWarps + 1) * 32;
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr int kMaxSmem = 227 * 1024;
constexpr int kRowDirect = 1 << 30;  // flag in a row's counter: one CTA attended the whole row and wrote bf16 itself
static_assert(kAttnIdDepth < kSchedQueue, "producers read kAttnIdDepth descriptors ahead of the one they issue");

template <int HPC>
struct AttnGeom {
  static constexpr int TL = kConsumerWarps / HPC;    // consumer warps per KV head
  static constexpr int TOK_BYTES = HPC * kRowBytes;   // one token's head-group row
  static constexpr int TOK_STRIDE = TOK_BYTES + 16;   // +16: ldmatrix rows hit distinct banks
  static constexpr int KV_BYTES = kPage * TOK_STRIDE;  // one tensor of one stage
  static constexpr int STAGE_BYTES = 2 * KV_BYTES;    // K then V
};

// consumer-only barrier (producer / scheduler warps never join)
__device__ __forceinline__ void consumer_sync() {
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory");
}

struct AttnSmemLayout {
  int ring_off, valid_off, bar_off, q_off, merge_off, rows_off, sched_off, total;
};

template <int HPC, int NT>
__host__ __device__ inline AttnSmemLayout attn_smem_layout(int stages, int gsz, int batch) {
  using G = AttnGeom<HPC>;
  AttnSmemLayout L;
  int off = 0;
  L.ring_off = off;   off += stages * G::STAGE_BYTES;
  L.q_off = off;      off += HPC * gsz * kRowBytes;
  L.merge_off = off;  off += (G::TL > 1) ? (kConsumerWarps - HPC) * 32 * NT * 36 * 4 : 0;
  L.bar_off = off;    off += 2 * stages * 8;
  L.valid_off = off;  off += ((stages * 4 + 15) / 16) * 16;
  L.rows_off = off;   off += (2 * batch * 4 + 15) / 16 * 16;  // clipped length, slot-table row per batch entry
  L.sched_off = off;  off += (int)((sizeof(SchedSmem) + 15) / 16 * 16);
  L.total = off;
  return L;
}

// contributor slots of one row: the CTAs that start on it, as many thieves, and some slack
__host__ __device__ inline int attn_row_slots(int grid, int rows) {
  const int cap = 2 * ((grid + rows - 1) / rows) + 24;
  return cap < grid ? cap : grid;
}

template <int HPC, int NT, bool WITH_SCORE>
__global__ void __launch_bounds__(kAttnThreads, 1) sparse_decode_attn_kernel(AttnParams p) {
  using G = AttnGeom<HPC>;
  constexpr int TL = G::TL;

  extern __shared__ __align__(128) unsigned char smem[];
  const int gsz = p.hq / p.hkv;
  const int HG = p.hkv / HPC;
  const int R = p.batch * HG;
  const int S = p.stages;
  const AttnSmemLayout L = attn_smem_layout<HPC, NT>(S, gsz, p.batch);
  unsigned char* ring = smem + L.ring_off;
  __nv_bfloat16* sQ = reinterpret_cast<__nv_bfloat16*>(smem + L.q_off);
  float* sMerge = reinterpret_cast<float*>(smem + L.merge_off);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + L.bar_off);
  uint64_t* empty_bar = full_bar + S;
  uint32_t* sValid = reinterpret_cast<uint32_t*>(smem + L.valid_off);
  int* sLen = reinterpret_cast<int*>(smem + L.rows_off);  // min(seq_len, max_len) of batch entry b
  int* sReq = sLen + p.batch;                             // slot-table row of batch entry b
  SchedSmem& sched = *reinterpret_cast<SchedSmem*>(smem + L.sched_off);
  __shared__ int s_flag, s_join;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned long long* dbg = p.dbg_times ? p.dbg_times + (size_t)blockIdx.x * 8 : nullptr;
  if (dbg && tid == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    dbg[0] = global_timer_ns();
    dbg[7] = smid;
    dbg[2] = dbg[4] = dbg[5] = dbg[6] = 0;
  }

  // ---- setup: barriers, per-entry lengths
  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 33);  // 32 async copy arrivals + the producer's release
      mbar_init(&empty_bar[s], HPC);
    }
    sched_init(sched, p.producers + kConsumerWarps);
    s_join = 0;
    mbar_fence_init();
  }
  for (int b = tid; b < p.batch; b += kAttnThreads) {
    sLen[b] = max(min(p.seq_lens[b], p.max_len), 0);
    sReq[b] = p.req_rows[b];
  }
  // rows without tokens produce zeros (reference: empty split -> 0 / -inf)
  for (int b = blockIdx.x; b < p.batch; b += gridDim.x) {
    if (p.seq_lens[b] <= 0 || p.max_len <= 0) {
      uint4* o4 = reinterpret_cast<uint4*>(reinterpret_cast<__nv_bfloat16*>(p.out) +
                                           (int64_t)b * p.out_stride_b);
      for (int i = tid; i < p.hq * 16; i += kAttnThreads) o4[i] = make_uint4(0, 0, 0, 0);
    }
  }
  __syncthreads();
  auto units_of_row = [&](int r) { return (sLen[r / HG] + kPage - 1) / kPage; };
  int* next_chunk = p.sched + 1;
  int* units_done = p.sched + 1 + R;
  int* n_contrib = p.sched + 1 + 2 * R;
  const int row_slots = attn_row_slots((int)gridDim.x, R);

  // ---- merge team: after their streaming roles the producer and scheduler warps merge the rows this
  // CTA is designated for (rows blockIdx, blockIdx + gridDim, ...): they poll the row's counter until
  // every unit is attended, then combine the contributors' partials (stage 2 of the reference).  The
  // consumers never wait for a merge or for a counter's round trip.
  constexpr int kTeamWarps = kProducerWarps + 1;
  auto team_sync = [&]() { asm volatile("bar.sync 2, %0;" ::"n"(kTeamWarps * 32) : "memory"); };
  auto merge_rows = [&]() {
    const int team_warp = warp - kConsumerWarps, ttid = tid - kConsumerThreads;
    const int heads_row = HPC * gsz;
    for (int row = blockIdx.x; row < R; row += gridDim.x) {
      const int n_row = units_of_row(row);
      if (n_row == 0) continue;
      if (ttid == 0) {
        const volatile int* ctr = units_done + row;
        int v = *ctr;
        if ((v & (kRowDirect - 1)) < n_row) {
          const uint64_t t0 = global_timer_ns();
          while (((v = *ctr) & (kRowDirect - 1)) < n_row) {
            __nanosleep(128);
            if (global_timer_ns() - t0 > 2000000000ull) __trap();
          }
        }
        s_flag = v;
      }
      team_sync();
      const bool direct = (s_flag & kRowDirect) != 0;
      team_sync();  // s_flag is rewritten for the next row
      if (direct) continue;
      const unsigned long long t_cmb = dbg ? global_timer_ns() : 0;
      const int b = row / HG, hg = row - b * HG;
      {
        // One warp per 4 q heads, lane = 4 output dims.  Lane c first fetches the state and lse of
        // contributor c (all contributors in parallel), then the partial outputs are read with
        // independent loads: the merge costs a couple of L2 round trips, not one per contributor.
        __threadfence();
        const int nc = min(__ldcg(n_contrib + row), row_slots);
        const int64_t base = (int64_t)row * row_slots;
        constexpr int HB = 4;  // heads a warp merges at once: their loads are all independent
        for (int h0 = team_warp * HB; h0 < heads_row; h0 += kTeamWarps * HB) {
          float m[HB], wsum[HB];
          float4 acc[HB];
#pragma unroll
          for (int x = 0; x < HB; ++x) m[x] = -INFINITY, wsum[x] = 0.f, acc[x] = make_float4(0.f, 0.f, 0.f, 0.f);
          for (int c0 = 0; c0 < nc; c0 += 32) {
            const int c = c0 + lane;
            const bool have = c < nc && __ldcg(p.slot_state + base + c) != 0;
            float lse[HB], wgt[HB];
#pragma unroll
            for (int x = 0; x < HB; ++x)
              lse[x] = (have && h0 + x < heads_row) ? __ldcg(p.ws_lse + (base + c) * heads_row + h0 + x)
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
              const float4* src = reinterpret_cast<const float4*>(
                                      p.ws_o + ((base + c0 + kc2) * heads_row + h0) * kHeadDim) + lane;
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
            }
          }
        }
      }
      team_sync();  // every warp has read the slot states
      for (int c = ttid; c < min(__ldcg(n_contrib + row), row_slots); c += kTeamWarps * 32)
        p.slot_state[(int64_t)row * row_slots + c] = 0;  // left zeroed
      if (dbg && ttid == 0) dbg[5] += global_timer_ns() - t_cmb;
    }
    // the CTA's scratch may be recycled once BOTH its consumers and its merge team are done
    if (ttid == 0 && atomicAdd(&s_join, 1) == 1) sched_finish(p.sched, R, R);
  };

  if (warp == kConsumerWarps + kProducerWarps) {
    // ============================================================== scheduler
    // a contributor slot of the row is reserved before the row is drained (aux of its descriptors)
    auto reserve = [&](int row) {
      int idx = 0;
      if (lane == 0) idx = atomicAdd(n_contrib + row, 1);
      idx = __shfl_sync(0xffffffffu, idx, 0);
      return idx < row_slots ? idx : -1;
    };
    sched_run(sched, next_chunk, R, min(R, (int)gridDim.x), kMinBatch, 4,
              min(kSchedQueue, (S + kChunkUnits - 1) / kChunkUnits + kAttnIdDepth + 1), units_of_row, reserve);
    merge_rows();
    return;
  }

  const __nv_bfloat16* kc = reinterpret_cast<const __nv_bfloat16*>(p.k_cache);
  const __nv_bfloat16* vc = reinterpret_cast<const __nv_bfloat16*>(p.v_cache);
  const int64_t slot_elems = (int64_t)p.hkv * kHeadDim;

  if (warp >= kConsumerWarps) {
    // ============================================================== producers
    // unit i (CTA-local running index) is fed by producer i % NP into stage i % S, NP | S: a stage is
    // always refilled by the same warp, so the empty-phase it waits for is never more than one ahead
    // of the barrier.  Lane l holds the slot of token unit_begin*16 + l of the chunk (2 units x 16
    // tokens), fetched kAttnIdDepth chunks ahead; each token row (HPC*256 contiguous bytes of K, same
    // of V) is copied by the whole warp with 16-byte LDGSTS (512 B per instruction), masked tokens
    // zero-filled.  Every lane's copies arrive on the stage's mbarrier when they land.
    const int pw = warp - kConsumerWarps;
    const int NP = p.producers;
    if (pw >= NP) {
      merge_rows();
      return;
    }
    ChunkDesc dq[kAttnIdDepth];
    int idq[kAttnIdDepth];
    int n_read = 0;
    bool ended = false;
    auto fetch = [&](ChunkDesc& d, int& id) {
      id = -1;
      if (ended) {
        d.row = -1;
        return;
      }
      d = sched_read(sched, n_read++);
      if (d.row < 0) {
        ended = true;
        return;
      }
      const int b = d.row / HG;
      const int t = d.unit_begin * kPage + lane;
      if (lane < d.units() * kPage && t < sLen[b])
        id = __ldg(p.slot_table + (int64_t)sReq[b] * p.slot_table_stride + t);
    };
#pragma unroll
    for (int t = 0; t < kAttnIdDepth; ++t) fetch(dq[t], idq[t]);
    constexpr int LPR = G::TOK_BYTES >= 512 ? 32 : G::TOK_BYTES / 16;  // lanes per token row
    constexpr int TPI = 32 / LPR;                                     // tokens per instruction
    constexpr int CPR = G::TOK_BYTES / (LPR * 16);                    // instructions per row
    const int sub = lane / LPR, chunk = lane % LPR;
    const int64_t slot_bytes = slot_elems * 2;
    // dq / idq: circular FIFO indexed at compile time (a register waiting for its table load is
    // never moved)
    bool done = false;
    while (!done) {
#pragma unroll
      for (int u = 0; u < kAttnIdDepth; ++u) {
        if (done) break;
        const ChunkDesc cur = dq[u];
        const int sa = idq[u];
        fetch(dq[u], idq[u]);
        if (cur.row < 0) {
          done = true;
          break;
        }
        const int my = (sa >= 0 && sa < p.num_slots) ? sa : -1;
        const uint32_t ball = __ballot_sync(0xffffffffu, my >= 0);
        const int hg = cur.row % HG;
        const char* kbase = reinterpret_cast<const char*>(kc) + hg * G::TOK_BYTES + chunk * 16;
        const char* vbase = reinterpret_cast<const char*>(vc) + hg * G::TOK_BYTES + chunk * 16;
        for (int j = 0; j < cur.units(); ++j) {
          const int i = cur.unit_base + j;
          if (i % NP != pw) continue;
          const int s = i % S;
          const uint32_t ph = (uint32_t)(i / S) & 1u;
          const uint32_t dst0 = smem_u32(ring + s * G::STAGE_BYTES) + chunk * 16;
          mbar_wait(&empty_bar[s], ph ^ 1u);
#pragma unroll
          for (int t0 = 0; t0 < kPage; t0 += TPI) {
            const int tok = t0 + sub;
            const int slot = __shfl_sync(0xffffffffu, my, j * kPage + tok);
            const int nbytes = slot >= 0 ? 16 : 0;
            const int64_t off = (int64_t)max(slot, 0) * slot_bytes;
            const uint32_t dst = dst0 + tok * G::TOK_STRIDE;
#pragma unroll
            for (int c = 0; c < CPR; ++c) {
              cp_async_16(dst + c * 512, kbase + off + c * 512, nbytes);
              cp_async_16(dst + G::KV_BYTES + c * 512, vbase + off + c * 512, nbytes);
            }
          }
          if (lane == 0) sValid[s] = (ball >> (j * kPage)) & 0xffffu;
          cp_async_mbar_arrive(&full_bar[s]);
          if (lane == 0) mbar_arrive(&full_bar[s]);  // release: publishes sValid
        }
      }
    }
    if (dbg && pw == 0 && lane == 0) dbg[4] = global_timer_ns();
    merge_rows();
    return;
  }

  // ================================================================ consumers
  const int g = lane >> 2, t = lane & 3;
  const int h_local = warp % HPC, tl = warp / HPC;
  const float scale_log2 = p.sm_scale * kLog2e;
  const int heads_row = HPC * gsz;

  // state of the current visit of a row
  uint32_t qb[NT][8][2];
  float m_run[NT][2], l_run[NT][2], o_acc[NT][8][4];
  int cur_row = -2, cur_slot = 0, visit_units = 0;  // -2: no row yet (-1 is the end descriptor)
  int total_units = 0, n_visits = 0;
  bool first_seen = false;

  // q rows of the row's head group -> smem -> B fragments (k = head dim, n = q head); fresh softmax state
  auto enter_row = [&](int row) {
    const int b = row / HG, hg = row - b * HG;
    const uint4* qsrc = reinterpret_cast<const uint4*>(
        reinterpret_cast<const __nv_bfloat16*>(p.q) + (int64_t)b * p.q_stride_b +
        (int64_t)hg * heads_row * kHeadDim);
    uint4* qdst = reinterpret_cast<uint4*>(sQ);
    for (int c = tid; c < heads_row * 16; c += kConsumerThreads) qdst[c] = qsrc[c];
    consumer_sync();
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
      m_run[nt][0] = m_run[nt][1] = -INFINITY;
      l_run[nt][0] = l_run[nt][1] = 0.f;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt)
#pragma unroll
        for (int e = 0; e < 4; ++e) o_acc[nt][dt][e] = 0.f;
    }
  };

  // one 16-token unit in ring stage s: tokens t0 .. t0+15 of batch entry b
  auto attend_unit = [&](int s, int b, int t0) {
    const uint32_t mask = sValid[s];
    const bool tv0 = (mask >> g) & 1u, tv1 = (mask >> (g + 8)) & 1u;
    if (mask == 0u) return;
    const uint32_t kbase = smem_u32(ring + s * G::STAGE_BYTES) + h_local * kRowBytes;
    const uint32_t vbase = kbase + G::KV_BYTES;
    float sc[NT][4];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
    {
      const uint32_t kaddr =
          kbase + ((lane & 7) + ((lane >> 3) & 1) * 8) * G::TOK_STRIDE + (lane >> 4) * 16;
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
        float* srow = p.attn_score + (int64_t)b * p.score_stride + t0;
        if (tv0) atomic_max_f32(srow + g, r0);
        if (tv1) atomic_max_f32(srow + g + 8, r1);
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
          vbase + ((lane & 7) + (lane >> 4) * 8) * G::TOK_STRIDE + ((lane >> 3) & 1) * 16;
#pragma unroll
      for (int dt = 0; dt < 8; ++dt) {
        uint32_t a[4];
        ldmatrix_x4_trans(a, vaddr + dt * 32);
#pragma unroll
        for (int nt = 0; nt < NT; ++nt) mma_16816(o_acc[nt][dt], a, pb[nt][0], pb[nt][1]);
      }
    }
  };

  // Leave a row: publish what this CTA attended.  A CTA that saw the whole row writes bf16 output;
  // otherwise one normalised fp32 partial + lse goes to the CTA's contributor slot of the row, the
  // units are added to the row's counter, and the CTA that completes the row merges all partials.
  auto leave_row = [&](int row, int slot, int units) {
    const unsigned long long t_epi = dbg ? global_timer_ns() : 0;
    const int b = row / HG, hg = row - b * HG;
    const int n_row = units_of_row(row);
    const bool complete = units == n_row;
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
      consumer_sync();
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
    const int64_t pslot = (int64_t)row * row_slots + slot;
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
              __nv_bfloat16* orow = reinterpret_cast<__nv_bfloat16*>(p.out) +
                                    (int64_t)b * p.out_stride_b +
                                    (int64_t)(hg * heads_row + head_local) * kHeadDim;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = __float2bfloat16_rn(o_acc[nt][dt][e] * inv);
                orow[dt * 16 + g + 8] = __float2bfloat16_rn(o_acc[nt][dt][2 + e] * inv);
              }
            } else {
              float* orow = p.ws_o + (pslot * heads_row + head_local) * kHeadDim;
#pragma unroll
              for (int dt = 0; dt < 8; ++dt) {
                orow[dt * 16 + g] = o_acc[nt][dt][e] * inv;
                orow[dt * 16 + g + 8] = o_acc[nt][dt][2 + e] * inv;
              }
              if (g == 0)
                p.ws_lse[pslot * heads_row + head_local] =
                    l > 0.f ? m_run[nt][e] * kLn2 + __logf(l) : -INFINITY;
            }
          }
        }
    }
    // publish: the row's counter is polled by the merge team of the row's designated CTA
    __threadfence();
    consumer_sync();
    if (tid == 0) {
      if (!complete) p.slot_state[pslot] = 1;  // this contributor slot holds a partial
      __threadfence();
      atomicAdd(units_done + row, complete ? (units | kRowDirect) : units);
    }
    consumer_sync();  // sQ / sMerge are reused by the next row
    if (dbg && tid == 0) dbg[6] += global_timer_ns() - t_epi;
  };

  // Unit i is consumed by the warps tl = i % TL; with TL | S a stage is always drained by the same
  // warps, so the phase a warp waits for is never more than one ahead of the barrier (parity waits
  // cannot tell a phase from the one two before it).
  for (int n = 0;; ++n) {
    const ChunkDesc d = sched_read(sched, n);
    if (d.row != cur_row) {
      if (cur_row >= 0) leave_row(cur_row, cur_slot, visit_units);
      if (d.row < 0) break;
      enter_row(d.row);
      cur_row = d.row;
      cur_slot = d.aux();
      visit_units = 0;
      ++n_visits;
    }
    visit_units += d.units();
    total_units += d.units();
    const int b = d.row / HG;
    for (int j = 0; j < d.units(); ++j) {
      const int i = d.unit_base + j;
      if (i % TL != tl) continue;
      const int s = i % S;
      mbar_wait(&full_bar[s], (uint32_t)(i / S) & 1u);
      if (dbg && tid == 0 && !first_seen) dbg[2] = global_timer_ns(), first_seen = true;
      attend_unit(s, b, (d.unit_begin + j) * kPage);
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[s]);
    }
  }
  if (tid == 0) {
    if (dbg) dbg[3] = global_timer_ns(), dbg[1] = (unsigned long long)total_units | ((unsigned long long)n_visits << 32);
    if (atomicAdd(&s_join, 1) == 1) sched_finish(p.sched, R, R);  // + the contributor counts
  }
}

template <int HPC, int NT, bool WITH_SCORE>
static int launch_variant(const AttnParams& p_in, cudaStream_t stream) {
  using G = AttnGeom<HPC>;
  AttnParams p = p_in;
  const int gsz = p.hq / p.hkv;
  const int rows = p.batch * (p.hkv / HPC);
  const int fixed = attn_smem_layout<HPC, NT>(0, gsz, p.batch).total + 256;  // + static smem
  int stages = (kMaxSmem - fixed) / (G::STAGE_BYTES + 20);
  if (stages > 24) stages = 24;
  // Unit i is consumed by the warps tl = i % TL and fed by producer i % NP; with TL | stages and
  // NP | stages a stage is always drained / refilled by the same warps (see the kernel).
  stages -= stages % G::TL;
  if (stages < G::TL) stages = 0;
  if (stages < 2) {
    set_error("sparse_decode_attn: batch=%d leaves no shared memory for the KV ring", p.batch);
    return 1;
  }
  p.stages = stages;
  p.dbg_times = debug_times_buffer();
  p.producers = 1;
  for (int d = kProducerWarps; d >= 1; --d)
    if (stages % d == 0) { p.producers = d; break; }
  const size_t smem = attn_smem_layout<HPC, NT>(stages, gsz, p.batch).total;
  auto kern = sparse_decode_attn_kernel<HPC, NT, WITH_SCORE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_error("sparse_decode_attn smem attr", e);
  const int64_t units_row = (p.max_len + kPage - 1) / kPage;
  const int64_t max_chunks = (int64_t)rows * ((units_row + kChunkUnits - 1) / kChunkUnits);
  int grid = max_ctas();
  if ((int64_t)grid > max_chunks) grid = (int)max_chunks;
  if (grid < 1) grid = 1;
  // the partial records (one per contributor slot of every row) must fit the caller's workspace
  const int heads_row = HPC * gsz;
  while (grid > 1 && ((int64_t)rows * attn_row_slots(grid, rows) * heads_row > p.rec_capacity ||
                      (int64_t)rows * attn_row_slots(grid, rows) > p.slot_capacity))
    --grid;
  if ((int64_t)rows * attn_row_slots(grid, rows) * heads_row > p.rec_capacity ||
      (int64_t)rows * attn_row_slots(grid, rows) > p.slot_capacity) {
    set_error("sparse_decode_attn: workspace too small (svb200_workspace_bytes gives the size)");
    return 1;
  }
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

int attn_auto_heads_per_cta(int hkv) {
  for (int c = 8; c > 1; c >>= 1)
    if (hkv % c == 0) return c;
  return 1;
}

int launch_sparse_decode_attn(const AttnParams& p, int heads_per_cta, cudaStream_t stream) {
  if (p.batch <= 0) return 0;
  const int hpc = heads_per_cta > 0 ? heads_per_cta : attn_auto_heads_per_cta(p.hkv);
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
