// Host-side declarations shared between the kernel translation units and api.cu.
#pragma once

#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace svb200 {

// error plumbing (api.cu)
void set_error(const char* fmt, ...);
int cuda_error(const char* what, cudaError_t e);
int check_launch(const char* what);
int sm_count();
unsigned long long* debug_times_buffer();
int max_ctas();  // persistent-grid size: sm_count() unless svb200_set_max_ctas lowered it
// experiment knobs (svb200_set_tuning); 0 = automatic
enum {
  kTuneScoreGroups = 0, kTuneAttnMinKB = 1, kTuneMinBatch = 2, kTuneShareDiv = 3, kTuneChunkUnits = 4,
  kTuneNoMath = 5, kTuneScoreQ = 6, kTunePdl = 7, kTuneAttnGranule = 8, kTuneAttnHpc = 9,
  kTuneAttnSplitPen = 10,  // attention: units a CTA is charged for a row boundary inside its range, + 1 (0: auto, 1: none)
  kTuneNoEvictFirst = 13,  // 1: the streams' bulk copies use the default L2 policy instead of evict-first
  kTunePollFirstNs = 14,   // one-launch layer: first sleep (ns) of a producer warp whose row is not selected yet
  kTuneLook = 15,          // score loader look-ahead in descriptors (1, 2, 4; 0: auto)
  kTuneTailChunks = 16,    // score: chunks of a row's tail that are claimed singly and posted unit by unit, + 1 (0: auto, 1: none)
  kTuneAttnRowCtas = 17,   // attention: most CTAs one row of work is cut over (0: auto)
  kTuneAttnStages = 11,    // attention: cap of the ring's stages (experiments; 0: as many as fit)
  kTuneCount = 24
};
int tuning(int which);
bool pdl_enabled();  // launches carry the programmatic-stream-serialization attribute

// Every kernel of the library is launched through this: with PDL enabled the launch may begin while its
// predecessor in the stream drains (the kernels call pdl_wait() before they touch global memory).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 Args&&... args) {
  static_assert(sizeof...(KArgs) == sizeof...(Args), "argument count");
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  unsigned n = 0;
  if (pdl_enabled()) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    n = 1;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// quest.cu
int launch_quest_page_minmax(const void* k_cache, void* bounds_layer, const int32_t* page_rec,
                             const int32_t* slot_mapping, int n_tokens, int hkv, int64_t num_slots,
                             int64_t num_chunks, cudaStream_t stream);
int launch_quest_decode_page_complete(const void* k_cache_all, void* bounds_all, const int32_t* page_rec,
                                      const int32_t* page_table, int64_t pt_stride,
                                      const int32_t* req_rows, const int32_t* seq_lens, int batch,
                                      int layers, int hkv, int64_t num_slots, int64_t num_pages,
                                      int64_t num_chunks, cudaStream_t stream);
int launch_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_chunks,
                            const int32_t* unit_table, int64_t ut_stride, const int32_t* req_rows,
                            const int32_t* context_lens, void* scores, int64_t score_stride,
                            int batch, int hq, int hkv, int n_cols, int32_t* sched, cudaStream_t stream);
int launch_quest_select_pack(const void* scores, int64_t score_stride, int n_cols,
                             const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
                             const int32_t* context_lens, int batch, int k, int32_t* packed_slots,
                             int64_t out_stride, int32_t* sparse_lens, int32_t* selected_pages,
                             const int32_t* dense_table, int64_t dense_stride, int token_budget,
                             int page_budget_base, int out_width, cudaStream_t stream);
// New K/V rows of the step (one token per batch row) stored by the scoring launch itself.
struct FusedStore {
  const void* key;
  int64_t k_stride;
  const void* value;
  int64_t v_stride;
  void* k_cache;
  void* v_cache;
  const int32_t* slot_mapping;
  int64_t num_slots;
  int refresh_bounds;  // rebuild the bounds record of a page the token completes
  const int32_t* page_rec;  // physical page -> bounds record (needed with refresh_bounds)
};

int launch_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds,
                                   int64_t num_chunks, const int32_t* unit_table, int64_t ut_stride,
                                   const int32_t* page_table, int64_t pt_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                   int n_cols, int k, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   const FusedStore* store, cudaStream_t stream);

// Peer access for the tensor-parallel epilogues (collective.cu, decode_attn.cu); mirrors svb200_comm_t.
constexpr int kMaxPeers = 8;
constexpr int kSyncEpoch = 8, kSyncDone = 9, kSyncInts = 16;  // int32 layout of a rank's sync block
struct PushArgs {
  int world, rank;             // world <= 1: no push
  __nv_bfloat16* out[kMaxPeers];  // gathered [B][world*Hq][128] bf16 buffer of every rank (own included)
  int64_t stride_b;            // elements between batch rows of the gathered buffers
  int32_t* sync[kMaxPeers];    // sync block of every rank: [r] = arrival flag written by rank r, [8] epoch, [9] finished CTAs
  // flag-in-data protocol (ll_words > 0): rank r's staging, two areas of ll_words 8-byte words {payload, epoch};
  // a row [src rank][b][local head] is 64 words
  unsigned long long* ll[kMaxPeers];
  int64_t ll_words;
};

// decode_attn.cu
struct AttnParams {
  const void* q;
  int64_t q_stride_b;
  const void* k_cache;
  const void* v_cache;
  int64_t num_slots;
  const int32_t* slot_table;
  int64_t slot_table_stride;
  const int32_t* req_rows;
  const int32_t* seq_lens;
  void* out;
  int64_t out_stride_b;
  float* attn_score;
  int64_t score_stride;
  int batch, hq, hkv, max_len;
  float* ws_o;       // [2*max_ctas][hq/head_groups][128] fp32 partial outputs
  float* ws_lse;     // [2*max_ctas][hq/head_groups]
  int32_t* tickets;  // [batch*head_groups] zeroed
  int max_ctas;      // CTAs the workspace was sized for
  int stages;        // filled by the launcher
  unsigned long long* dbg_times;  // experiments: per-CTA timestamps (svb200_debug_times), else null
  int producers;     // producer warps in use (divides stages); filled by the launcher
  float sm_scale;
  int no_math;       // experiments: consumers release the stages without the contractions
  int granule;       // copy granules of several token rows where the slots allow it
  int grid_cap;      // CTAs the attention work is cut into at most (0: the whole grid); set by a fused launcher
  int evict_first;   // the K/V stream's copies are tagged evict-first in L2; launcher
  int split_pen;     // units a CTA is charged for a row boundary inside its range (see the kernel's cut); launcher
  PushArgs push;     // tensor-parallel push epilogue (world <= 1: none)
};
int attn_auto_heads_per_cta(int hkv, int batch, int max_len);
int attn_row_ctas(int hpc, int max_len);
// launch configuration of the attention phase of a one-launch layer (quest.cu): ring stages inside `region_bytes` of
// shared memory, producer warps, CTAs the work is cut into (<= grid, <= the workspace's CTAs), knobs
int attn_configure_fused(AttnParams& p, int heads_per_cta, int region_bytes, int grid);
int launch_quest_decode_layer(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_chunks,
                              const int32_t* unit_table, int64_t ut_stride, const int32_t* page_table,
                              int64_t pt_stride, const int32_t* req_rows, const int32_t* context_lens,
                              void* scores, int64_t score_stride, int batch, int hq, int hkv, int n_cols, int k,
                              int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens,
                              int32_t* selected_pages, int32_t* sched, const FusedStore* store,
                              const AttnParams& attn, cudaStream_t stream);
int launch_sparse_decode_attn(const AttnParams& p, int heads_per_cta, cudaStream_t stream);

// h2o.cu
int launch_h2o_normalize_accumulate(float* score, int64_t score_stride, int batch, int width,
                                    float scale, float* cumulative, int64_t cum_stride,
                                    const int32_t* cum_rows, const int32_t* kv_lens, float refill,
                                    cudaStream_t stream);
int launch_h2o_select_compact(float* cumulative, const int64_t* cum_rows, int32_t* slot_tables,
                              const int64_t* slot_rows, int32_t* free_stack,
                              const int64_t* free_rows, int32_t* keep_out, int num_rows,
                              const int32_t* kv_lens, int kv_len, int budget, int recent, cudaStream_t stream);

int launch_h2o_prepare_decode(const int32_t* meta, const int32_t* free_stack, int64_t free_stride, int32_t* slot_tables,
                              int64_t layer_stride, int64_t row_stride, int32_t* layer_slot_mapping, int64_t lsm_stride,
                              int32_t* context_lens, int32_t* req_indices, int32_t* slot_mapping, int32_t* kv_lens,
                              int32_t* rows_out, int batch, int layers, cudaStream_t stream);
struct PrefillScoreHostArgs {
  const void* q; int64_t q_stride;
  const void* k_cache; int64_t num_slots;
  const int32_t* slot_table; int64_t table_stride;
  const int32_t* req_rows; const int32_t* start_loc; const int32_t* context_lens;
  const int32_t* prompt_cache_lens; const int32_t* score_starts; const int32_t* score_ends;
  float* score; int64_t score_stride;
  float* cumulative; int64_t cum_stride; const int32_t* cum_rows;
  int batch, hq, hkv, max_query_len, max_context_len, candidate_start, num_recent;
  float sm_scale;
};
int launch_h2o_prefill_score(const PrefillScoreHostArgs& h, cudaStream_t stream);

// prefill_attn.cu
struct PrefillAttnHostArgs {
  const void* q; int64_t q_stride;
  const void* k_cache; const void* v_cache; int64_t num_slots;
  const int32_t* slot_table; int64_t table_stride;
  const int32_t* req_rows; const int32_t* start_loc; const int32_t* context_lens; const int32_t* prompt_cache_lens;
  void* out; int64_t out_stride;
  int batch, hq, hkv, max_chunk_len;
  float sm_scale;
};
int launch_prefill_attn(const PrefillAttnHostArgs& h, cudaStream_t stream);

// collective.cu
struct AllReduceArgs {
  int world, rank;
  const __nv_bfloat16* in[kMaxPeers];  // this call's input of every rank (own included), n elements each
  int32_t* sync[kMaxPeers];
  __nv_bfloat16* out;                  // local result
  int64_t n;                           // elements, a multiple of 8
};
int launch_allreduce_sum_bf16(const AllReduceArgs& a, cudaStream_t stream);

// store_kv.cu
int launch_store_kv_minmax(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                           void* k_cache, void* v_cache, const int32_t* slot_mapping, int n_tokens,
                           int hkv, int64_t num_slots, void* bounds_layer, const int32_t* page_rec,
                           int64_t num_chunks, cudaStream_t stream);

}  // namespace svb200
