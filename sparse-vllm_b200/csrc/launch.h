// Host-side declarations shared between the kernel translation units and api.cu.
#pragma once

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
enum { kTuneScoreGroups = 0, kTuneAttnMinKB = 1, kTuneMinBatch = 2, kTuneShareDiv = 3, kTuneScoreCopy = 4, kTuneScoreSplit = 5, kTuneScoreQ = 6, kTuneCount = 8 };
int tuning(int which);

// quest.cu
int launch_quest_page_minmax(const void* k_cache, void* bounds_layer, const int32_t* slot_mapping,
                             int n_tokens, int hkv, int64_t num_slots, cudaStream_t stream);
int launch_quest_decode_page_complete(const void* k_cache_all, void* bounds_all,
                                      const int32_t* page_table, int64_t pt_stride,
                                      const int32_t* req_rows, const int32_t* seq_lens, int batch,
                                      int layers, int hkv, int64_t num_slots, int64_t num_pages,
                                      cudaStream_t stream);
int launch_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds, int64_t num_pages,
                            const int32_t* page_table, int64_t pt_stride, const int32_t* req_rows,
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
};

int launch_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds,
                                   int64_t num_pages, const int32_t* page_table, int64_t pt_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int batch, int hq, int hkv,
                                   int n_cols, int k, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   const FusedStore* store, cudaStream_t stream);

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
};
int attn_auto_heads_per_cta(int hkv);
int launch_sparse_decode_attn(const AttnParams& p, int heads_per_cta, cudaStream_t stream);

// h2o.cu
int launch_h2o_normalize_accumulate(float* score, int64_t score_stride, int batch, int width,
                                    float scale, float* cumulative, int64_t cum_stride,
                                    const int32_t* kv_lens, cudaStream_t stream);
int launch_h2o_select_compact(float* cumulative, const int64_t* cum_rows, int32_t* slot_tables,
                              const int64_t* slot_rows, int32_t* free_stack,
                              const int64_t* free_rows, int32_t* keep_out, int num_rows, int kv_len,
                              int budget, int recent, cudaStream_t stream);

// store_kv.cu
int launch_store_kv_minmax(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                           void* k_cache, void* v_cache, const int32_t* slot_mapping, int n_tokens,
                           int hkv, int64_t num_slots, void* bounds_layer, cudaStream_t stream);

}  // namespace svb200
