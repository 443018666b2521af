/*
 * svb200.h — C ABI of the B200 (sm_100a) provider for Sparse-vLLM's decode-time
 * sparse-attention hot path (Quest page selection + sparse paged GQA decode
 * attention, H2O scored decode).
 *
 * Conventions
 *   - Every pointer is a DEVICE pointer unless the parameter name ends in _host.
 *   - All tensors are bf16 (uint16 storage) / fp32 / int32 as stated; head_dim is
 *     fixed at 128, quest page size at 16.
 *   - Every call is asynchronous on `stream` (a cudaStream_t passed as void*), does
 *     no allocation and no host synchronisation, and is CUDA-graph capturable.
 *   - Return value: 0 on success; SVB200_E* on validation failure; 1000 + cudaError_t
 *     when a launch failed.  svb200_last_error() returns a thread-local message.
 *   - No ownership transfer: buffers belong to the caller and must outlive the call.
 *
 * "Reference" below is CURRENTF/Sparse-vLLM, paths relative to its src/sparsevllm/.
 */
#ifndef SVB200_H_
#define SVB200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SVB200_ABI_VERSION 6

enum {
  SVB200_OK = 0,
  SVB200_EINVAL = 1,      /* bad shape / alignment / null pointer            */
  SVB200_EUNSUPPORTED = 2 /* head_dim != 128, page != 16, group size > 16 …  */
};

#define SVB200_HEAD_DIM 128
#define SVB200_PAGE 16

/* ABI / build introspection. */
int svb200_abi_version(void);
const char* svb200_last_error(void);
/* Number of SMs of the current device (cached); used by hosts to size split-K. */
int svb200_sm_count(void);
/*
 * Cap the grid of the persistent kernels (score/select and attention) at n CTAs;
 * n <= 0 restores one CTA per SM.  Returns the previous cap (0 = none).  Process-wide
 * tuning / test knob: fewer CTAs means more page tiles per CTA and rows cut at
 * different places; results do not depend on it.
 */
int svb200_set_max_ctas(int32_t n);
/*
 * Experiment knob `which` := value; 0 restores the automatic choice.  Returns the previous value,
 * -1 for an unknown knob.  Results never depend on it (except 5, which is an experiment that skips the math).
 * Score kernel: 0 = row groups of the scheduler, 2 = chunks per claim at least, 3 = divisor of the guided
 * claim size, 4 = units per claimed chunk, 6 = 1: stage q once / 2: stage q+ and q-.
 * Attention kernel: 1 = KB of K/V a CTA should stream at least (grid sizing of small launches; default 256),
 * 8 = 2: one bulk copy per token row even where a granule of rows is contiguous, 9 = KV heads staged together by
 * a CTA when the caller passes heads_per_cta = 0 (1, 2, 4, 8; default: the largest of 4, 2, 1 that divides Hkv).
 * Both: 5 = 1 (score) / 2 (attention): stream the bytes but skip the tensor-core contractions (results
 * are garbage; measures what the copy pipeline alone sustains), 7 = programmatic dependent launch
 * (1 = on, 2 = off): a launch may start while its predecessor in the stream drains.
 */
int svb200_set_tuning(int32_t which, int32_t value);
/*
 * int32 entries of the scheduler scratch the persistent score kernels need for `rows` batch rows
 * (1 + 2*rows: finished-CTA count, next chunk per row, scored units per row).  Zero-initialise it
 * once; the last CTA of every launch leaves it zeroed.  One scratch per stream: two launches that
 * may run concurrently must not share it.
 */
int64_t svb200_sched_scratch_ints(int32_t rows);
/*
 * Experiments only: when device_buffer != NULL the attention kernel writes 8 uint64 per CTA
 * (globaltimer ns at entry, after setup, first data, consumers done, producer done, -, -, SM id).
 */
void svb200_debug_times(void* device_buffer);

/*
 * Bytes of scratch svb200_sparse_decode_attn needs: one int32 arrival ticket per
 * (batch row, KV head) followed by two fp32 partial records (output + lse) per SM.
 * The buffer must be zero-initialised once; kernels leave the tickets zeroed.
 * Takes the place of the reference's mid_o / mid_o_logexpsum split-K workspace
 * (layers/attention.py:15-51), whose size grows with batch * context / 256; this one
 * depends on the SM count only.
 */
int64_t svb200_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads);

/* ------------------------------------------------------------------------- *
 * Quest page bounds ("metadata").  Layout owned by this provider (csrc/bounds_layout.cuh):
 * a *bounds chunk* holds the bounds of 16 consecutive LOGICAL pages of one request row,
 *   bounds_layer[num_chunks][hs][16][2 (max,min)][Hkv/hs][128] bf16
 * (hs = smallest divisor of Hkv leaving <= 4 heads; 16-byte segments of every 256-byte row are
 *  XOR-swizzled by page & 7), so the scoring scan reads one contiguous 8-32 KB piece per ring stage
 * with no page-table indirection.  The reference keeps two planes metadata_cache[2][L][P][Hkv][D]
 * indexed by PHYSICAL page and gathers them per row (engine/cache_manager/quest.py:96-105,
 * 1870-1881).  Two int32 maps, maintained by the cache manager when it allocates pages:
 *   unit_table[rows][stride]: bounds chunk of (row, logical page / 16), -1 = none
 *   page_rec[num_pages]     : physical page -> chunk * 16 + (logical page % 16), -1 = unassigned
 * ------------------------------------------------------------------------- */

/*
 * Recompute min/max of every page COMPLETED by the tokens in slot_mapping
 * (slot % 16 == 15, slot >= 0) from the K cache of one layer.
 * Replaces QuestCacheManager.on_kv_stored / _on_kv_stored_prefill_capture
 * (engine/cache_manager/quest.py:1608-1716): K rows were already stored, so
 * full pages and pages completed across chunks are the same scan.
 *   k_cache      [num_slots][Hkv][128] bf16 (one layer)
 *   bounds_layer num_chunks bounds chunks (one layer); page_rec [num_slots/16] int32
 *   slot_mapping [n_tokens] int32 (-1 = padding)
 */
int svb200_quest_page_minmax(const void* k_cache, void* bounds_layer, int64_t num_chunks,
                             const int32_t* page_rec, const int32_t* slot_mapping,
                             int32_t n_tokens, int32_t num_kv_heads, int64_t num_slots,
                             void* stream);

/*
 * Decode-time page completion for ALL layers in one launch: for every batch row
 * whose length after this step is a positive multiple of 16, rebuild the bounds
 * of page (len/16 - 1) in each layer.
 * Replaces QuestCacheManager.on_forward_end (quest.py:1719-1771, a Python loop
 * over layers with three launches each).
 *   k_cache_all  [L][num_slots][Hkv][128] bf16, bounds_all [L][num_chunks] bounds chunks
 *   page_table   [rows][page_table_stride] int32, req_rows [batch] int32,
 *   seq_lens     [batch] int32 (length INCLUDING the token just stored)
 */
int svb200_quest_decode_page_complete(const void* k_cache_all, void* bounds_all, int64_t num_chunks,
                                      const int32_t* page_rec,
                                      const int32_t* page_table,
                                      int64_t page_table_stride,
                                      const int32_t* req_rows,
                                      const int32_t* seq_lens, int32_t batch,
                                      int32_t num_layers, int32_t num_kv_heads,
                                      int64_t num_slots, int64_t num_pages,
                                      void* stream);

/*
 * Page scores for one layer: score[b][p] for p in [0, n_cols) where
 *   score = amax over all local q heads of
 *           bf16( bf16(sum_d q+[d]*max[p][d]) + bf16(sum_d q-[d]*min[p][d]) )
 * (fp32 accumulation, three bf16 roundings — exactly the rounding structure of
 *  QuestCacheManager._score_pages_batched, quest.py:1773-1802), columns
 * p >= ceil(ctx/16) - 1 are -inf (quest.py:1883-1885).  The gather of
 * quest.py:1870-1881 does not exist: a row's bounds are contiguous chunks.
 *   q            [batch][Hq][128] bf16, row stride q_stride_b elements
 *   bounds_layer num_chunks bounds chunks; units whose unit_table entry is outside
 *                [0, num_chunks) score -inf
 *   unit_table   [rows][unit_table_stride] int32
 *   req_rows     [batch] int32 row of each batch entry in unit_table / page_table
 *   context_lens [batch] int32
 *   scores       [batch][score_stride] bf16 out, score_stride >= n_cols
 *   sched_scratch svb200_sched_scratch_ints(batch) zeroed int32 (left zeroed)
 * Work is claimed dynamically (~64 KB chunks from per-row counters), so SMs the memory
 * system serves faster score more pages; the result does not depend on who scored what.
 */
int svb200_quest_page_score(const void* q, int64_t q_stride_b,
                            const void* bounds_layer, int64_t num_chunks,
                            const int32_t* unit_table,
                            int64_t unit_table_stride, const int32_t* req_rows,
                            const int32_t* context_lens, void* scores,
                            int64_t score_stride, int32_t batch, int32_t num_q_heads,
                            int32_t num_kv_heads, int32_t n_cols, int32_t* sched_scratch,
                            void* stream);

/*
 * Top-k page selection + slot packing for one layer (quest.py:1887-1913):
 * picks the prev_budget highest-scoring columns (ties: lowest page index first),
 * appends the last page, expands page slots to token slots.
 *   packed_slots [batch][out_stride] int32, out_stride >= (prev_budget+1)*16
 *   sparse_lens  [batch] int32 = prev_budget*16 + (ctx - (ceil(ctx/16)-1)*16)
 *   selected_pages [batch][prev_budget+1] int32 or NULL (logical page ids, ascending
 *                  then the last page; for tests / debugging)
 * If dense_token_table != NULL (batch is not long-text, quest.py:1866-1868,
 * 1903-1911) rows with ctx <= token_budget or ceil(ctx/16) <= page_budget_base take
 * the first out_width entries of their dense row instead and keep ctx as length;
 * out_width = max_keep >= (prev_budget+1)*16 columns are written.
 */
int svb200_quest_select_pack(const void* scores, int64_t score_stride,
                             int32_t n_cols, const int32_t* page_table,
                             int64_t page_table_stride, const int32_t* req_rows,
                             const int32_t* context_lens, int32_t batch,
                             int32_t prev_budget, int32_t* packed_slots,
                             int64_t out_stride, int32_t* sparse_lens,
                             int32_t* selected_pages,
                             const int32_t* dense_token_table,
                             int64_t dense_table_stride, int32_t token_budget,
                             int32_t page_budget_base, int32_t out_width,
                             void* stream);

/*
 * Fused score + select + pack: same results as the two calls above with one
 * launch (the CTA that scores a row's last pages performs its selection while the
 * others go on with the next rows).
 *   tickets  svb200_sched_scratch_ints(batch) zero-initialised int32 scratch (left zeroed).
 */
int svb200_quest_score_select_pack(
    const void* q, int64_t q_stride_b, const void* bounds_layer, int64_t num_chunks,
    const int32_t* unit_table, int64_t unit_table_stride,
    const int32_t* page_table, int64_t page_table_stride, const int32_t* req_rows,
    const int32_t* context_lens, void* scores, int64_t score_stride, int32_t batch,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols, int32_t prev_budget,
    int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens,
    int32_t* selected_pages, int32_t* tickets, void* stream);

/*
 * No counterpart in the reference: added for callers that cannot allocate
 * device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 * svb200_quest_score_select_pack with the scheduler scratch (`tickets`) inside the library: a file-scope
 * `static __device__` array of 1 + 3 * SVB200_STATIC_MAX_BATCH int32 (~12 KB, resident on every GPU that loads
 * the library until the process exits), whose device address the library takes once per device with
 * cudaGetSymbolAddress.  The call allocates nothing (no cudaMalloc / cudaFree, no allocating library routine);
 * it clears the part of the array it uses on `stream`, then launches the same kernel.  Inputs, outputs and results
 * are those of svb200_quest_score_select_pack.
 *   batch > SVB200_STATIC_MAX_BATCH -> SVB200_EINVAL with a message (use the caller-owned-scratch entry point).
 * Every call on a device shares the one array: calls that overlap on the GPU (different streams, or graphs replayed
 * concurrently) race on it; serialise them or use svb200_quest_score_select_pack with one scratch per stream.
 */
#define SVB200_STATIC_MAX_BATCH 1024
int svb200_quest_score_select_pack_static(
    const void* q, int64_t q_stride_b, const void* bounds_layer, int64_t num_chunks,
    const int32_t* unit_table, int64_t unit_table_stride,
    const int32_t* page_table, int64_t page_table_stride, const int32_t* req_rows,
    const int32_t* context_lens, void* scores, int64_t score_stride, int32_t batch,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols, int32_t prev_budget,
    int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens,
    int32_t* selected_pages, void* stream);

/*
 * svb200_quest_score_select_pack that also performs the decode step's KV store
 * (svb200_store_kv_minmax for batch tokens, token b = batch row b) inside the same
 * launch: warps that wait for the first row to be scored write the rows.  Legal because
 * a decode token always lands in its row's LAST page, which the launch never scores
 * (columns [0, pages-1)), and the attention launch that reads the rows comes after.
 * Replaces store_kvcache (kernels/triton/store_kvcache.py:10-71) + the decode part
 * of on_forward_end (quest.py:1719-1771) + build_decode_view (quest.py:1805-1913) for
 * one layer.
 *   refresh_bounds  != 0: rebuild the bounds record of a page the token completes
 *                   (through page_rec); bounds_layer is written.
 * With n_cols <= 0 nothing is scored but the rows are still stored.
 */
int svb200_quest_store_score_select_pack(
    const void* key, int64_t k_stride, const void* value, int64_t v_stride,
    void* k_cache, void* v_cache, const int32_t* slot_mapping, int64_t num_slots,
    int32_t refresh_bounds, const int32_t* page_rec,
    const void* q, int64_t q_stride_b, void* bounds_layer, int64_t num_chunks,
    const int32_t* unit_table, int64_t unit_table_stride,
    const int32_t* page_table, int64_t page_table_stride, const int32_t* req_rows,
    const int32_t* context_lens, void* scores, int64_t score_stride, int32_t batch,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols, int32_t prev_budget,
    int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens,
    int32_t* selected_pages, int32_t* tickets, void* stream);

/*
 * Sparse paged GQA decode attention: stream-K over 16-token pages with a fused
 * combine (one launch).
 * Replaces gqa_flash_decode_stage1(+_with_score 2D) and flash_decode_stage2
 * (kernels/triton/gqa_flash_decoding_stage1.py:6-121,211-295,
 *  kernels/triton/flash_decoding_stage2.py:8-46) behind
 * TritonAttentionBackend.run_decode (layers/attention_backend.py:217-349).
 *   q          [batch][Hq][128] bf16 (row stride q_stride_b elements)
 *   k_cache, v_cache [num_slots][Hkv][128] bf16, identical layout
 *   slot_table [rows][slot_table_stride] int32 token slots; entries outside
 *              [0, num_slots) are treated as masked tokens
 *   max_len    upper bound of seq_lens (<= slot_table_stride); longer rows are clipped
 *   req_rows   [batch] int32, seq_lens [batch] int32 (tokens visible per row)
 *   out        [batch][Hq][128] bf16 (row stride out_stride_b elements); rows with
 *              seq_len <= 0 are zero-filled
 *   attn_score NULL (Quest) or fp32 [batch][score_stride] pre-filled by the caller
 *              (reference: -1e20); entry n is max-updated with the raw, unscaled
 *              max over local q heads of q.k[n]  (H2O / SnapKV 2D score).
 *   heads_per_cta 0 = auto, else 1/2/4/8 KV heads staged together by one CTA.
 *   workspace  zero-initialised scratch of svb200_workspace_bytes() bytes (a smaller
 *              buffer limits the number of CTAs).
 */
int svb200_sparse_decode_attn(const void* q, int64_t q_stride_b, const void* k_cache,
                              const void* v_cache, int64_t num_slots,
                              const int32_t* slot_table, int64_t slot_table_stride,
                              int32_t max_len, const int32_t* req_rows,
                              const int32_t* seq_lens, void* out, int64_t out_stride_b,
                              float* attn_score, int64_t score_stride, int32_t batch,
                              int32_t num_q_heads, int32_t num_kv_heads,
                              int32_t heads_per_cta, void* workspace,
                              int64_t workspace_bytes, float sm_scale, void* stream);

/*
 * H2O layer-end normalisation (+ optional fused accumulation).
 *   score [batch][score_stride] fp32, in place: score = softmax(score * scale) over
 *   the first `width` columns (engine/sparse_controller.py:761-767).
 *   If cumulative != NULL: for batch entry b with kv_len = kv_lens[b] and row c = cum_rows[b] (b when
 *   cum_rows is NULL: consecutive rows):
 *     cumulative[c][i] += score[b][i] for i < kv_len-1; cumulative[c][kv_len-1] = score
 *   (H2OCacheManager.update_decode_attention_scores_all_layers, h2o.py:958-1038,
 *    without the stack/clone of per-row tensors).
 *   refill != 0 (with cumulative): the score row is left holding `refill` (the reference's -1e20) instead of
 *   the normalised values, i.e. ready for the next step's max-updates; the reference refills the whole
 *   [L,B,W] buffer with a launch of its own every step (sparse_controller.py:545-594).
 */
int svb200_h2o_normalize_accumulate(float* score, int64_t score_stride, int32_t batch,
                                    int32_t width, float scale, float* cumulative,
                                    int64_t cum_stride, const int32_t* cum_rows,
                                    const int32_t* kv_lens, float refill, void* stream);

/*
 * Device side of H2OCacheManager.prepare_decode_static (h2o.py:257-476) in one launch: meta = int32
 * [rows[B], new_lens[B], free_top] (host-packed, one pinned copy).  For every (layer l, batch entry b) the slot
 * free_stack[l][free_top - B + b] is appended to the layer's slot row (slot_tables[l][row][new_len - 1]) and
 * written to layer_slot_mapping[l][b]; layer 0 also fills the static step buffers context_lens / req_indices /
 * slot_mapping / kv_lens / rows_out [B].
 */
int svb200_h2o_prepare_decode(const int32_t* meta, const int32_t* free_stack, int64_t free_stride,
                              int32_t* slot_tables, int64_t layer_stride, int64_t row_stride,
                              int32_t* layer_slot_mapping, int64_t lsm_stride, int32_t* context_lens,
                              int32_t* req_indices, int32_t* slot_mapping, int32_t* kv_lens,
                              int32_t* rows_out, int32_t batch, int32_t num_layers, void* stream);

/*
 * Dense causal prefill attention of a chunk over its paged row (cached prefix + the chunk, whose K/V rows are
 * already stored): query m of row b sits at physical position m + prompt_cache_lens[b] and attends to kv
 * positions <= its own through the slot table; logits q.k * sm_scale, fp32 online softmax, P cast to bf16
 * before PV.  Replaces context_attention_fwd (kernels/triton/context_flashattention_nopad.py:11-79, 242-302).
 *   q, out [T][Hq][128] bf16 (token strides q_stride / out_stride), start_loc[b] = first chunk token of row b
 *   max_chunk_len = longest chunk of the batch (sizes the grid)
 */
int svb200_prefill_attn(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache,
                        int64_t num_slots, const int32_t* slot_table, int64_t slot_table_stride,
                        const int32_t* req_rows, const int32_t* start_loc, const int32_t* context_lens,
                        const int32_t* prompt_cache_lens, void* out, int64_t out_stride, int32_t batch,
                        int32_t num_q_heads, int32_t num_kv_heads, int32_t max_chunk_len, float sm_scale,
                        void* stream);

/*
 * Prefill window scoring (H2O; the SnapKV-style candidate range is supported too): for batch row b the
 * queries at physical positions [score_starts[b], score_ends[b]) — the last <= 128 tokens of the prefill
 * chunk — attend causally to the row's kv positions [candidate_start, max(candidate_start, ctx - num_recent));
 *   score[b][kv] = max over q heads of (sum over those queries of softmax_kv(q.k * sm_scale)[q][kv]) / n_queries
 * Replaces prefill_score_fwd in probability mode (kernels/triton/prefill_score.py:70-428: partial-stats,
 * reduce and final kernels) as driven by H2OCacheManager.collect_prefill_attention_score (h2o.py:778-894).
 *   q [T][Hq][128] bf16 chunk tokens of all rows (token stride q_stride), start_loc[b] = first token of row b
 *   k_cache [num_slots][Hkv][128] bf16 (the chunk's K rows are already stored), slot_table [rows][stride]
 *   context_lens / prompt_cache_lens [B]: physical row length with / without the chunk
 *   score [B][score_stride] fp32: overwritten (zero outside the candidate range)
 * If cumulative != NULL the step scores are accumulated like H2OCacheManager._accumulate_score (h2o.py:611-626):
 *   cumulative[cum_rows[b]][i] = (i < prompt_cache_len ? cumulative[..][i] : 0) + n_queries * score[b][i], i < ctx
 */
int svb200_h2o_prefill_score(const void* q, int64_t q_stride, const void* k_cache, int64_t num_slots,
                             const int32_t* slot_table, int64_t slot_table_stride,
                             const int32_t* req_rows, const int32_t* start_loc,
                             const int32_t* context_lens, const int32_t* prompt_cache_lens,
                             const int32_t* score_starts, const int32_t* score_ends, float* score,
                             int64_t score_stride, float* cumulative, int64_t cum_stride,
                             const int32_t* cum_rows, int32_t batch, int32_t num_q_heads,
                             int32_t num_kv_heads, int32_t max_query_len, int32_t max_context_len,
                             int32_t candidate_start, int32_t num_recent_tokens, float sm_scale,
                             void* stream);

/*
 * H2O burst eviction for `num_rows` (layer,row) pairs of length kv_lens[i] (all kv_len when kv_lens is
 * NULL; kv_len is then also the longest row: it sizes the shared-memory row cache; entries no longer than
 * `budget` are left alone):
 * heavy hitters = the (budget - recent) largest cumulative scores among
 * [0, kv_len - recent) (stable: lower index wins ties), plus the `recent` suffix,
 * in ascending order (H2OCacheManager.select_h2o_indices_batch, h2o.py:519-563);
 * then index-only compaction of the slot rows, scores and the per-layer free
 * stack (SnapKVCacheManager.free_part_slots_batch_layers, snapkv.py:1681-1803).
 *   cum_rows   int64 [num_rows] element offsets of each row in `cumulative`
 *   slot_rows  int64 [num_rows] element offsets of each row in `slot_tables`
 *   free_rows  int64 [num_rows] element offset in `free_stack` where this row's
 *              dropped slots start (its kv_len - budget entries), or NULL
 *   keep_out   int32 [num_rows][budget] ascending kept indices (may be NULL)
 */
int svb200_h2o_select_compact(float* cumulative, const int64_t* cum_rows,
                              int32_t* slot_tables, const int64_t* slot_rows,
                              int32_t* free_stack, const int64_t* free_rows,
                              int32_t* keep_out, int32_t num_rows, const int32_t* kv_lens,
                              int32_t kv_len, int32_t budget, int32_t recent, void* stream);

/*
 * Store the new K/V rows of one layer and (quest_bounds_layer != NULL) refresh the
 * bounds of any page the stored token completes.
 * Replaces store_kvcache (kernels/triton/store_kvcache.py:10-71) and, for decode,
 * the deferred refresh of quest.py:1719-1771.
 *   key, value [n_tokens][Hkv][128] bf16 with token strides k_stride/v_stride (elements)
 *   slot_mapping [n_tokens] int32, -1 = skip
 */
int svb200_store_kv_minmax(const void* key, int64_t k_stride, const void* value,
                           int64_t v_stride, void* k_cache, void* v_cache,
                           const int32_t* slot_mapping, int32_t n_tokens,
                           int32_t num_kv_heads, int64_t num_slots,
                           void* quest_bounds_layer, const int32_t* page_rec,
                           int64_t num_chunks, void* stream);

/* ------------------------------------------------------------------------- *
 * Tensor parallelism over NVLink peer memory (one process per GPU).
 * A communicator is plain data: every rank maps every rank's data buffer and sync block
 * (svb200_symm_alloc / svb200_ipc_open below, or any other peer-mapped memory).
 *   peer_buf[r]   data buffer of rank r (same size on every rank), 16-byte aligned
 *   peer_sync[r]  sync block of rank r: int32[16], zero-initialised; [q] = arrival flag written by
 *                 rank q, [8] = epoch, [9] = finished CTAs (owner only)
 * Every rank must issue the same sequence of collective launches on a communicator.  A rank can run at
 * most one collective ahead of its peers, so callers alternate two regions of peer_buf.
 * ------------------------------------------------------------------------- */
#define SVB200_MAX_PEERS 8
typedef struct {
  int32_t world, rank;
  void* peer_buf[SVB200_MAX_PEERS];
  int32_t* peer_sync[SVB200_MAX_PEERS];
  /* Optional staging of the push epilogue's flag-in-data protocol (ll_bytes = 0: none; the epilogue then orders its
   * peer stores with system-scope fences and per-rank flags).  peer_ll[r]: ll_bytes of zero-initialised,
   * 16-byte aligned memory of rank r, mapped by every rank.  A launch whose gathered tensor has G bytes needs
   * ll_bytes >= 4 * G: every 4 payload bytes travel as one 8-byte word {payload, epoch}, which a 64-bit store
   * makes visible atomically, so the receiver needs no fence and the sender no flag; two areas alternate. */
  void* peer_ll[SVB200_MAX_PEERS];
  int64_t ll_bytes;
} svb200_comm_t;

/*
 * svb200_sparse_decode_attn whose epilogue also gathers the head shards: every final bf16 row
 * [b][h_local] is stored to rank r's peer_buf[r] + gather_offset_bytes at
 * [b][rank*Hq + h_local][128] (row stride gather_stride_b elements >= world*Hq*128) for every r, and the
 * launch completes only after all ranks' rows have arrived here — the [B][Hq_total][128] tensor o_proj reads.
 * Replaces the per-layer NCCL collective after attention (the reference all-reduces after a row-parallel
 * o_proj, layers/linear.py:581-589; its fused providers, operators/all_reduce.py:115-409, are H100-only).
 */
int svb200_sparse_decode_attn_tp(const void* q, int64_t q_stride_b, const void* k_cache,
                                 const void* v_cache, int64_t num_slots,
                                 const int32_t* slot_table, int64_t slot_table_stride,
                                 int32_t max_len, const int32_t* req_rows,
                                 const int32_t* seq_lens, void* out, int64_t out_stride_b,
                                 float* attn_score, int64_t score_stride, int32_t batch,
                                 int32_t num_q_heads, int32_t num_kv_heads,
                                 int32_t heads_per_cta, void* workspace,
                                 int64_t workspace_bytes, float sm_scale,
                                 const svb200_comm_t* comm, int64_t gather_offset_bytes,
                                 int64_t gather_stride_b, void* stream);

/*
 * One launch per sparse decode layer: svb200_quest_store_score_select_pack (key == NULL: no store) followed, inside
 * the same launch, by svb200_sparse_decode_attn(_tp) over the slots it packed.  Replaces, for a long-text batch,
 * save_rope_kv_if_needed + _build_decode_view_static (engine/cache_manager/quest.py:1773-1913) + run_decode
 * (layers/attention_backend.py:217-349) of one layer.  A CTA whose scoring work is done sets the attention up while
 * the last rows are still being selected, and reads a row's packed slots as soon as that row's selection is
 * published: the launch gap, the attention's set-up and its ramp overlap the selection tail.
 * Results are those of the two launches (same kernels' code).  Needs n_cols > 0 and at most 8 q heads per KV head
 * (else SVB200_EUNSUPPORTED: use the separate entry points).
 *   sched_scratch svb200_sched_scratch_ints(batch) zero-initialised int32 (left zeroed)
 *   workspace     as for svb200_sparse_decode_attn
 *   comm          NULL, or the peers of the tensor-parallel push epilogue (svb200_sparse_decode_attn_tp)
 */
int svb200_quest_decode_layer(
    const void* key, int64_t k_stride, const void* value, int64_t v_stride, const int32_t* slot_mapping,
    int32_t refresh_bounds, const int32_t* page_rec, void* k_cache, void* v_cache, int64_t num_slots,
    const void* q, int64_t q_stride_b, void* bounds_layer, int64_t num_chunks,
    const int32_t* unit_table, int64_t unit_table_stride,
    const int32_t* page_table, int64_t page_table_stride, const int32_t* req_rows,
    const int32_t* context_lens, void* scores, int64_t score_stride, int32_t batch,
    int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols, int32_t prev_budget,
    int32_t* packed_slots, int64_t out_stride, int32_t* sparse_lens, int32_t* selected_pages,
    int32_t* sched_scratch, void* out, int64_t out_stride_b, void* workspace, int64_t workspace_bytes,
    float sm_scale, const svb200_comm_t* comm, int64_t gather_offset_bytes, int64_t gather_stride_b,
    void* stream);


/*
 * One-shot SUM all-reduce of n_elems bf16 (a multiple of 8): rank r's input is
 * peer_buf[r] + in_offset_bytes; every rank reads all inputs over NVLink, accumulates in fp32 in rank
 * order (bit-identical results on all ranks) and writes bf16 to its local `out`.
 * Role of ParallelContext.tp_all_reduce after RowParallelLinear (distributed/parallel_context.py:87-116,
 * layers/linear.py:581-589; provider family operators/all_reduce.py).
 */
int svb200_allreduce_sum_bf16(const svb200_comm_t* comm, int64_t in_offset_bytes, int64_t n_elems,
                              void* out, void* stream);

/* cudaMalloc'ed, zeroed device memory + its 64-byte CUDA IPC handle (host buffers). */
int svb200_symm_alloc(int64_t bytes, void** ptr_host, void* handle64_host);
int svb200_symm_free(void* ptr);
/* Map a peer's allocation from its IPC handle (enables peer access lazily); unmap it. */
int svb200_ipc_open(const void* handle64_host, void** ptr_host);
int svb200_ipc_close(void* ptr);

#ifdef __cplusplus
}
#endif
#endif /* SVB200_H_ */
