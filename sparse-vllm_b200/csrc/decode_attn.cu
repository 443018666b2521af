// Sparse paged GQA decode attention for sm_100a: the stand-alone kernel and its launchers (device code and the
// description of the decomposition: decode_attn.cuh).
#include <cmath>
#include "decode_attn.cuh"

namespace svb200 {

template <int HPC, int NT, bool WITH_SCORE>
__global__ void __launch_bounds__(kAttnThreads, 1) sparse_decode_attn_kernel(AttnParams p) {
  extern __shared__ __align__(128) unsigned char smem[];
  const int tid = threadIdx.x;
  unsigned long long* dbg = p.dbg_times ? p.dbg_times + (size_t)blockIdx.x * 8 : nullptr;
  if (dbg && tid == 0) {
    uint32_t smid;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    dbg[0] = global_timer_ns();
    dbg[7] = smid;
    dbg[5] = dbg[6] = 0;
  }
  if (tid < kConsumerThreads)
    attn_setup<HPC, NT>(p, smem, [&](int b) { return min(p.seq_lens[b], p.max_len); });
  else
    pdl_wait();
  __syncthreads();
  pdl_launch_dependents();
  if (dbg && tid == 0) dbg[1] = global_timer_ns();
  attn_main<HPC, NT, WITH_SCORE, 1, false>(p, smem, [](int) {}, dbg);
}

// Launch configuration shared by the stand-alone kernel and the attention phase of a one-launch layer: ring stages
// that fit `avail` bytes of shared memory, producer warps, CTAs the work is cut into (<= grid_in), knobs.
// Most CTAs one row of work (HPC KV heads of a batch entry, <= max_len tokens) is cut over.  The row's last
// contributor merges one partial per CTA that had units of it (~0.65 us each) while a CTA streams ~45 GB/s, so
// time(G) = bytes / (G * 45 GB/s) + 0.65 us * G is lowest at G = sqrt(row KB / 29).  Attention launch, us, at
// 16 / 12 / 10 / 8 CTAs per row: bs 1 ctx 32K budget 2048 (4 MB rows) 19.4 / 17.5 / 18.8 / 20.1, bs 4 21.0 / 18.3 /
// 18.9 / 20.2 (GPU call 49); dense 128K rows (268 MB) get 95, i.e. every SM; a dense 32K layer at bs 1 (two rows of
// 34 MB, 74 CTAs each before) went from 61.6 to 48.3 us (calls 57-58).
int attn_row_ctas(int hpc, int max_len) {
  if (tuning(kTuneAttnRowCtas) > 0) return tuning(kTuneAttnRowCtas);
  const double row_kb = (double)((max_len + kPage - 1) / kPage) * (hpc * 2 * kPage * kHeadDim * 2 / 1024);
  const int n = (int)std::ceil(std::sqrt(row_kb / 29.0));
  return n < 4 ? 4 : n;
}

template <int HPC, int NT>
static int attn_configure(AttnParams& p, int avail, int grid_in, int* grid_out) {
  using G = AttnGeom<HPC>;
  const int gsz = p.hq / p.hkv;
  const int rows = p.batch * (p.hkv / HPC);
  const int fixed = attn_smem_layout<HPC, NT>(0, gsz, rows, p.batch).total;
  int stages = (avail - fixed) / (G::STAGE_BYTES + 20);
  if (stages > 24) stages = 24;
  if (tuning(kTuneAttnStages) > 0 && stages > tuning(kTuneAttnStages)) stages = tuning(kTuneAttnStages);
  // Unit j is consumed by the warps tl = j % TL; with TL | stages a stage is always drained by the
  // same warps, so the phase a warp waits for is never more than one ahead of the barrier (parity
  // waits cannot tell a phase from the one two before it).
  stages -= stages % G::TL;
  if (stages < G::TL) stages = 0;
  if (stages < 2) {
    set_error("sparse_decode_attn: batch*head_groups=%d rows leave no shared memory for the KV ring", rows);
    return 1;
  }
  p.stages = stages;
  p.producers = 1;
  for (int d = kProducerWarps; d >= 1; --d)
    if (stages % d == 0) { p.producers = d; break; }
  const int64_t max_units = (int64_t)rows * ((p.max_len + kPage - 1) / kPage);
  int grid = grid_in;
  if (grid > p.max_ctas) grid = p.max_ctas;
  if ((int64_t)grid > max_units) grid = (int)max_units;
  // Small launches (one short row: cfg2) are latency-bound: every CTA pays the same start-up and the
  // last contributor of a row merges one partial per CTA, so a CTA should stream >= ~256 KB of K/V
  // (bs 1, 2064 tokens, 8 KV heads: 148 CTAs 36 us, 32 CTAs 17 us).  max_len bounds the work from above.
  const int64_t unit_kb = (int64_t)HPC * 2 * kPage * kHeadDim * 2 / 1024;  // K + V rows of HPC heads, 16 tokens
  {
    int kb = tuning(kTuneAttnMinKB) > 0 ? tuning(kTuneAttnMinKB) : 256;
    const int64_t min_units = (kb + unit_kb - 1) / unit_kb;
    const int64_t want = (max_units + min_units - 1) / min_units;
    if ((int64_t)grid > want) grid = (int)want;
  }
  // ... and a row of work is cut over no more CTAs than its bytes pay for (attn_row_ctas)
  {
    const int per_row = attn_row_ctas(HPC, p.max_len);
    if ((int64_t)grid > (int64_t)rows * per_row) grid = rows * per_row;
  }
  if (grid < 1) grid = 1;
  *grid_out = grid;
  p.grid_cap = 0;
  p.no_math = tuning(kTuneNoMath) == 2;
  p.granule = tuning(kTuneAttnGranule) != 2;
  p.evict_first = tuning(kTuneNoEvictFirst) != 1;
  // ~6 us of streaming at ~44 GB/s per SM, in units (the kernel caps it at half a range)
  p.split_pen = tuning(kTuneAttnSplitPen) > 0 ? tuning(kTuneAttnSplitPen) - 1
                                               : (int)((6 * 44000 + unit_kb * 512) / (unit_kb * 1024));
  return 0;
}

int attn_configure_fused(AttnParams& p, int heads_per_cta, int region_bytes, int grid) {
  int cap = 0, rc = 1;
  switch (heads_per_cta) {
    case 1: rc = attn_configure<1, 1>(p, region_bytes, grid, &cap); break;
    case 2: rc = attn_configure<2, 1>(p, region_bytes, grid, &cap); break;
    case 4: rc = attn_configure<4, 1>(p, region_bytes, grid, &cap); break;
    default: set_error("quest_decode_layer: %d KV heads per CTA not supported", heads_per_cta);
  }
  if (rc) return rc;
  p.grid_cap = cap;
  return 0;
}

template <int HPC, int NT, bool WITH_SCORE>
static int launch_variant(const AttnParams& p_in, cudaStream_t stream) {
  AttnParams p = p_in;
  int grid = 0;
  if (int rc = attn_configure<HPC, NT>(p, kMaxSmem - 256 /* static smem */, max_ctas(), &grid)) return rc;
  p.dbg_times = debug_times_buffer();
  const size_t smem = attn_smem_layout<HPC, NT>(p.stages, p.hq / p.hkv, p.batch * (p.hkv / HPC), p.batch).total;
  auto kern = sparse_decode_attn_kernel<HPC, NT, WITH_SCORE>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) return cuda_error("sparse_decode_attn smem attr", e);
  e = launch_kernel(kern, dim3(grid), dim3(kAttnThreads), smem, stream, p);
  if (e != cudaSuccess) return cuda_error("sparse_decode_attn launch", e);
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

int attn_auto_heads_per_cta(int hkv, int batch, int max_len) {
  // 4 heads = 1 KB token rows and six ring stages.  8 heads (knob 9 = 1): three stages of 64 KB, but with all of a
  // rank's heads staged together the rows of a page are contiguous and a copy moves a granule of them.
  if (tuning(kTuneAttnHpc) > 0 && hkv % tuning(kTuneAttnHpc) == 0) return tuning(kTuneAttnHpc);
  // Small batches: with 4 heads per CTA there are too few rows of work to occupy the SMs (bs 1, 8 KV heads: 2 rows);
  // 2 heads per CTA doubles the rows.  Attention launch, us, 4 -> 2 heads per CTA at ctx 128K budget 4096: bs 1
  // 25.7 -> 20.5, bs 2 25.7 -> 20.6, bs 4 26.6 -> 23.6, bs 8 34.5 -> 33.9 (stays at 4); ctx 32K budget 2048: bs 1-4
  // unchanged within 0.3 us (GPU call 51).
  if (hkv % 4 == 0 && (int64_t)batch * (hkv / 4) * attn_row_ctas(4, max_len) < max_ctas()) return 2;
  for (int c = 4; c > 1; c >>= 1)
    if (hkv % c == 0) return c;
  return 1;
}

int launch_sparse_decode_attn(const AttnParams& p, int heads_per_cta, cudaStream_t stream) {
  if (p.batch <= 0) return 0;
  const int hpc = heads_per_cta > 0 ? heads_per_cta : attn_auto_heads_per_cta(p.hkv, p.batch, p.max_len);
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
