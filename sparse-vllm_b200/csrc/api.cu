// extern "C" entry points of libsvb200.so (see include/svb200.h): argument
// validation, error reporting, dispatch to the kernel launchers.  No torch types.
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "../../include/svb200.h"
#include "launch.h"

namespace svb200 {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_error(const char* what, cudaError_t e) {
  set_error("%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
  return 1000 + (int)e;
}
int check_launch(const char* what) {
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return cuda_error(what, e);
  return 0;
}
int sm_count() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      cached = 148;
  }
  return cached;
}

#ifndef SVB200_PDL_DEFAULT
#define SVB200_PDL_DEFAULT 0
#endif
static unsigned long long* g_dbg_times = nullptr;
unsigned long long* debug_times_buffer() { return g_dbg_times; }
static int g_tuning[kTuneCount] = {0};
int tuning(int which) { return (which >= 0 && which < kTuneCount) ? g_tuning[which] : 0; }
// programmatic dependent launch (knob 7: 1 = on, 2 = off, 0 = the build's default)
bool pdl_enabled() { return g_tuning[kTunePdl] != 2 && (g_tuning[kTunePdl] == 1 || SVB200_PDL_DEFAULT); }
static int g_max_ctas = 0;
int max_ctas() {
  const int n = sm_count();
  return (g_max_ctas > 0 && g_max_ctas < n) ? g_max_ctas : n;
}

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

#define SV_REQUIRE(cond, ...)    \
  do {                           \
    if (!(cond)) {               \
      set_error(__VA_ARGS__);    \
      return SVB200_EINVAL;      \
    }                            \
  } while (0)
#define SV_SUPPORTED(cond, ...)  \
  do {                           \
    if (!(cond)) {               \
      set_error(__VA_ARGS__);    \
      return SVB200_EUNSUPPORTED;\
    }                            \
  } while (0)

static int check_heads(const char* who, int hq, int hkv) {
  SV_REQUIRE(hq > 0 && hkv > 0 && hq % hkv == 0, "%s: num_q_heads=%d must be a positive multiple of num_kv_heads=%d", who, hq, hkv);
  SV_SUPPORTED(hq / hkv <= 16, "%s: GQA group size %d > 16 is not supported", who, hq / hkv);
  return 0;
}

}  // namespace svb200

using namespace svb200;

extern "C" {

int svb200_abi_version(void) { return SVB200_ABI_VERSION; }
const char* svb200_last_error(void) { return g_err; }
int svb200_sm_count(void) { return sm_count(); }
void svb200_debug_times(void* device_buffer) { g_dbg_times = (unsigned long long*)device_buffer; }
int svb200_set_max_ctas(int32_t n) {
  const int prev = g_max_ctas;
  g_max_ctas = n > 0 ? n : 0;
  return prev;
}

int svb200_set_tuning(int32_t which, int32_t value) {
  if (which < 0 || which >= kTuneCount) return -1;
  const int prev = g_tuning[which];
  g_tuning[which] = value;
  return prev;
}
int64_t svb200_sched_scratch_ints(int32_t rows) { return rows < 0 ? -1 : 1 + 3 * (int64_t)rows; }

static int64_t ws_ticket_bytes(int64_t batch, int64_t hkv) {
  return ((batch * hkv * (int64_t)sizeof(int32_t) + 255) / 256) * 256;
}
static int64_t ws_cta_bytes(int64_t hq) { return 2 * hq * (SVB200_HEAD_DIM + 4) * (int64_t)sizeof(float); }

int64_t svb200_workspace_bytes(int32_t batch, int32_t num_q_heads, int32_t num_kv_heads) {
  if (batch < 0 || num_q_heads <= 0 || num_kv_heads <= 0) return -1;
  return ws_ticket_bytes(batch, num_kv_heads) + (int64_t)sm_count() * ws_cta_bytes(num_q_heads);
}

int svb200_quest_page_minmax(const void* k_cache, void* bounds_layer, int64_t num_chunks,
                             const int32_t* page_rec, const int32_t* slot_mapping,
                             int32_t n_tokens, int32_t num_kv_heads, int64_t num_slots,
                             void* stream) {
  SV_REQUIRE(n_tokens >= 0, "quest_page_minmax: n_tokens=%d < 0", n_tokens);
  if (n_tokens == 0) return 0;
  SV_REQUIRE(k_cache && bounds_layer && slot_mapping && page_rec, "quest_page_minmax: null pointer");
  SV_REQUIRE(num_chunks > 0 && num_chunks <= (INT32_MAX >> 4), "quest_page_minmax: num_chunks=%lld out of range", (long long)num_chunks);
  SV_REQUIRE(aligned16(k_cache) && aligned16(bounds_layer), "quest_page_minmax: buffers must be 16-byte aligned");
  SV_REQUIRE(num_kv_heads > 0 && num_slots > 0 && num_slots % SVB200_PAGE == 0,
             "quest_page_minmax: num_kv_heads=%d num_slots=%lld (must be a positive multiple of 16)",
             num_kv_heads, (long long)num_slots);
  return launch_quest_page_minmax(k_cache, bounds_layer, page_rec, slot_mapping, n_tokens, num_kv_heads,
                                  num_slots, num_chunks, (cudaStream_t)stream);
}

int svb200_quest_decode_page_complete(const void* k_cache_all, void* bounds_all, int64_t num_chunks,
                                      const int32_t* page_rec,
                                      const int32_t* page_table, int64_t page_table_stride,
                                      const int32_t* req_rows, const int32_t* seq_lens,
                                      int32_t batch, int32_t num_layers, int32_t num_kv_heads,
                                      int64_t num_slots, int64_t num_pages, void* stream) {
  SV_REQUIRE(batch >= 0 && num_layers >= 0, "quest_decode_page_complete: negative batch/layers");
  if (batch == 0 || num_layers == 0) return 0;
  SV_REQUIRE(k_cache_all && bounds_all && page_table && req_rows && seq_lens && page_rec,
             "quest_decode_page_complete: null pointer");
  SV_REQUIRE(num_chunks > 0 && num_chunks <= (INT32_MAX >> 4), "quest_decode_page_complete: num_chunks=%lld out of range", (long long)num_chunks);
  SV_REQUIRE(aligned16(k_cache_all) && aligned16(bounds_all), "quest_decode_page_complete: buffers must be 16-byte aligned");
  SV_REQUIRE(num_kv_heads > 0 && num_slots == num_pages * SVB200_PAGE && page_table_stride > 0,
             "quest_decode_page_complete: num_slots=%lld must equal 16*num_pages=%lld",
             (long long)num_slots, (long long)num_pages);
  SV_REQUIRE(num_layers <= 65535, "quest_decode_page_complete: num_layers=%d > 65535", num_layers);
  return launch_quest_decode_page_complete(k_cache_all, bounds_all, page_rec, page_table, page_table_stride,
                                           req_rows, seq_lens, batch, num_layers, num_kv_heads,
                                           num_slots, num_pages, num_chunks, (cudaStream_t)stream);
}

static int check_score_args(const char* who, const void* q, int64_t q_stride_b, const void* bounds,
                            int64_t num_chunks, const int32_t* unit_table, int64_t ut_stride, const int32_t* req_rows,
                            const int32_t* ctx, const void* scores, int64_t score_stride,
                            int batch, int hq, int hkv, int n_cols) {
  SV_REQUIRE(q && bounds && unit_table && req_rows && ctx && scores, "%s: null pointer", who);
  SV_REQUIRE(aligned16(q) && aligned16(bounds) && aligned16(scores), "%s: q/bounds/scores must be 16-byte aligned", who);
  if (int rc = check_heads(who, hq, hkv)) return rc;
  SV_REQUIRE(q_stride_b >= (int64_t)hq * SVB200_HEAD_DIM && q_stride_b % 8 == 0,
             "%s: q_stride_b=%lld too small or not a multiple of 8", who, (long long)q_stride_b);
  SV_REQUIRE(n_cols > 0 && ut_stride * SVB200_PAGE >= n_cols, "%s: unit table stride %lld must cover n_cols=%d pages",
             who, (long long)ut_stride, n_cols);
  // the selection scans tie counts in 16-bit lanes of one word
  SV_SUPPORTED(n_cols < 32768, "%s: n_cols=%d >= 32768 pages (contexts of 512K tokens and more) is not supported", who, n_cols);
  SV_REQUIRE(score_stride >= ((n_cols + 7) / 8) * 8 && score_stride % 8 == 0,
             "%s: score_stride=%lld must be a multiple of 8 covering n_cols=%d", who,
             (long long)score_stride, n_cols);
  SV_REQUIRE(num_chunks > 0, "%s: num_chunks=%lld must be positive", who, (long long)num_chunks);
  SV_REQUIRE(batch <= 8192, "%s: batch=%d > 8192", who, batch);
  return 0;
}

int svb200_quest_page_score(const void* q, int64_t q_stride_b, const void* bounds_layer,
                            int64_t num_chunks, const int32_t* unit_table, int64_t unit_table_stride,
                            const int32_t* req_rows, const int32_t* context_lens, void* scores,
                            int64_t score_stride, int32_t batch, int32_t num_q_heads,
                            int32_t num_kv_heads, int32_t n_cols, int32_t* sched_scratch, void* stream) {
  SV_REQUIRE(batch >= 0, "quest_page_score: batch=%d < 0", batch);
  if (batch == 0) return 0;
  SV_REQUIRE(sched_scratch != nullptr, "quest_page_score: sched_scratch is null");
  if (int rc = check_score_args("quest_page_score", q, q_stride_b, bounds_layer, num_chunks, unit_table,
                                unit_table_stride, req_rows, context_lens, scores, score_stride,
                                batch, num_q_heads, num_kv_heads, n_cols))
    return rc;
  return launch_quest_page_score(q, q_stride_b, bounds_layer, num_chunks, unit_table, unit_table_stride,
                                 req_rows, context_lens, scores, score_stride, batch, num_q_heads,
                                 num_kv_heads, n_cols, sched_scratch, (cudaStream_t)stream);
}

static int check_select_args(const char* who, int n_cols, int prev_budget, const int32_t* packed,
                             int64_t out_stride, const int32_t* lens) {
  SV_REQUIRE(packed && lens, "%s: null output pointer", who);
  SV_REQUIRE(aligned16(packed) && out_stride % 4 == 0, "%s: packed_slots must be 16-byte aligned with a stride multiple of 4", who);
  SV_REQUIRE(prev_budget >= 1 && prev_budget <= n_cols, "%s: prev_budget=%d must be in [1, n_cols=%d]", who, prev_budget, n_cols);
  SV_SUPPORTED(prev_budget <= 2047, "%s: prev_budget=%d > 2047 pages is not supported", who, prev_budget);
  SV_REQUIRE(out_stride >= (int64_t)(prev_budget + 1) * SVB200_PAGE, "%s: out_stride=%lld < (prev_budget+1)*16", who, (long long)out_stride);
  return 0;
}

int svb200_quest_select_pack(const void* scores, int64_t score_stride, int32_t n_cols,
                             const int32_t* page_table, int64_t page_table_stride,
                             const int32_t* req_rows, const int32_t* context_lens, int32_t batch,
                             int32_t prev_budget, int32_t* packed_slots, int64_t out_stride,
                             int32_t* sparse_lens, int32_t* selected_pages,
                             const int32_t* dense_token_table, int64_t dense_table_stride,
                             int32_t token_budget, int32_t page_budget_base, int32_t out_width,
                             void* stream) {
  SV_REQUIRE(batch >= 0, "quest_select_pack: batch=%d < 0", batch);
  if (batch == 0) return 0;
  SV_REQUIRE(scores && page_table && req_rows && context_lens, "quest_select_pack: null pointer");
  SV_REQUIRE(aligned16(scores) && score_stride % 8 == 0 && score_stride >= ((n_cols + 7) / 8) * 8,
             "quest_select_pack: scores must be 16-byte aligned with stride multiple of 8 covering n_cols");
  SV_REQUIRE(n_cols > 0 && page_table_stride >= n_cols + 1, "quest_select_pack: page table stride must cover n_cols+1 pages");
  SV_SUPPORTED(n_cols < 32768, "quest_select_pack: n_cols=%d >= 32768 pages is not supported", n_cols);
  if (int rc = check_select_args("quest_select_pack", n_cols, prev_budget, packed_slots, out_stride, sparse_lens)) return rc;
  if (dense_token_table != nullptr) {
    SV_REQUIRE(out_width >= (prev_budget + 1) * SVB200_PAGE && out_stride >= out_width && dense_table_stride >= out_width,
               "quest_select_pack: out_width=%d must be >= (prev_budget+1)*16 and fit both strides", out_width);
  }
  return launch_quest_select_pack(scores, score_stride, n_cols, page_table, page_table_stride,
                                  req_rows, context_lens, batch, prev_budget, packed_slots,
                                  out_stride, sparse_lens, selected_pages, dense_token_table,
                                  dense_table_stride, token_budget, page_budget_base, out_width,
                                  (cudaStream_t)stream);
}

int svb200_quest_score_select_pack(const void* q, int64_t q_stride_b, const void* bounds_layer,
                                   int64_t num_chunks, const int32_t* unit_table, int64_t unit_table_stride,
                                   const int32_t* page_table, int64_t page_table_stride,
                                   const int32_t* req_rows, const int32_t* context_lens,
                                   void* scores, int64_t score_stride, int32_t batch,
                                   int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols,
                                   int32_t prev_budget, int32_t* packed_slots, int64_t out_stride,
                                   int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                   void* stream) {
  SV_REQUIRE(batch >= 0, "quest_score_select_pack: batch=%d < 0", batch);
  if (batch == 0) return 0;
  if (int rc = check_score_args("quest_score_select_pack", q, q_stride_b, bounds_layer, num_chunks, unit_table,
                                unit_table_stride, req_rows, context_lens, scores, score_stride,
                                batch, num_q_heads, num_kv_heads, n_cols))
    return rc;
  if (int rc = check_select_args("quest_score_select_pack", n_cols, prev_budget, packed_slots, out_stride, sparse_lens)) return rc;
  SV_REQUIRE(tickets != nullptr, "quest_score_select_pack: tickets scratch is null");
  SV_REQUIRE(page_table && page_table_stride >= n_cols + 1, "quest_score_select_pack: page table (stride %lld) must cover n_cols+1=%d pages",
             (long long)page_table_stride, n_cols + 1);
  return launch_quest_score_select_pack(q, q_stride_b, bounds_layer, num_chunks, unit_table, unit_table_stride,
                                        page_table, page_table_stride,
                                        req_rows, context_lens, scores, score_stride, batch,
                                        num_q_heads, num_kv_heads, n_cols, prev_budget,
                                        packed_slots, out_stride, sparse_lens, selected_pages,
                                        tickets, nullptr, (cudaStream_t)stream);
}

// ---- svb200_quest_score_select_pack_static: the scheduler scratch lives in the library (no allocation anywhere)
// The array's name is a host-side stand-in in host code, never a device address: the address a kernel may use is
// taken once per device with cudaGetSymbolAddress (with that device current) and cached.
static __device__ __align__(256) int32_t g_static_sched_scratch[1 + 3 * SVB200_STATIC_MAX_BATCH + 15];
constexpr int kStaticMaxDevices = 64;
static int32_t* g_static_scratch_ptr[kStaticMaxDevices] = {nullptr};

static int static_sched_scratch(const char* who, int32_t** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_error("cudaGetDevice", e);
  if (dev < 0 || dev >= kStaticMaxDevices) {
    set_error("%s: device ordinal %d beyond the %d this library keeps static scratch addresses for", who, dev, kStaticMaxDevices);
    return SVB200_EUNSUPPORTED;
  }
  // (two threads racing here both obtain the same address: the store is idempotent)
  if (g_static_scratch_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_sched_scratch);
    if (e != cudaSuccess) return cuda_error("cudaGetSymbolAddress(static scheduler scratch)", e);
    if (ptr == nullptr) {
      set_error("%s: cudaGetSymbolAddress returned a null device address", who);
      return SVB200_EINVAL;
    }
    g_static_scratch_ptr[dev] = static_cast<int32_t*>(ptr);
  }
  *out = g_static_scratch_ptr[dev];
  return 0;
}

int svb200_quest_score_select_pack_static(const void* q, int64_t q_stride_b, const void* bounds_layer,
                                          int64_t num_chunks, const int32_t* unit_table, int64_t unit_table_stride,
                                          const int32_t* page_table, int64_t page_table_stride,
                                          const int32_t* req_rows, const int32_t* context_lens,
                                          void* scores, int64_t score_stride, int32_t batch,
                                          int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols,
                                          int32_t prev_budget, int32_t* packed_slots, int64_t out_stride,
                                          int32_t* sparse_lens, int32_t* selected_pages, void* stream) {
  const char* who = "quest_score_select_pack_static";
  SV_REQUIRE(batch >= 0, "%s: batch=%d < 0", who, batch);
  if (batch == 0) return 0;
  SV_REQUIRE(batch <= SVB200_STATIC_MAX_BATCH,
             "%s: batch=%d exceeds SVB200_STATIC_MAX_BATCH=%d rows the library's static scratch holds "
             "(use svb200_quest_score_select_pack with a caller-owned scratch)", who, batch, SVB200_STATIC_MAX_BATCH);
  int32_t* scratch = nullptr;
  if (int rc = static_sched_scratch(who, &scratch)) return rc;
  // the array keeps what the last call left (a launch leaves it zeroed, an aborted one may not): clear what this
  // call uses, on the caller's stream (a memset node under graph capture; nothing is allocated)
  const cudaError_t e = cudaMemsetAsync(scratch, 0, sizeof(int32_t) * (size_t)svb200_sched_scratch_ints(batch), (cudaStream_t)stream);
  if (e != cudaSuccess) return cuda_error("quest_score_select_pack_static: clearing the static scratch", e);
  return svb200_quest_score_select_pack(q, q_stride_b, bounds_layer, num_chunks, unit_table, unit_table_stride,
                                        page_table, page_table_stride, req_rows, context_lens, scores, score_stride,
                                        batch, num_q_heads, num_kv_heads, n_cols, prev_budget, packed_slots,
                                        out_stride, sparse_lens, selected_pages, scratch, stream);
}

int svb200_quest_store_score_select_pack(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                                         void* k_cache, void* v_cache, const int32_t* slot_mapping,
                                         int64_t num_slots, int32_t refresh_bounds, const int32_t* page_rec,
                                         const void* q, int64_t q_stride_b, void* bounds_layer,
                                         int64_t num_chunks, const int32_t* unit_table, int64_t unit_table_stride,
                                         const int32_t* page_table, int64_t page_table_stride,
                                         const int32_t* req_rows, const int32_t* context_lens,
                                         void* scores, int64_t score_stride, int32_t batch,
                                         int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols,
                                         int32_t prev_budget, int32_t* packed_slots, int64_t out_stride,
                                         int32_t* sparse_lens, int32_t* selected_pages, int32_t* tickets,
                                         void* stream) {
  const char* who = "quest_store_score_select_pack";
  SV_REQUIRE(batch >= 0, "%s: batch=%d < 0", who, batch);
  if (batch == 0) return 0;
  if (int rc = check_score_args(who, q, q_stride_b, bounds_layer, num_chunks, unit_table,
                                unit_table_stride, req_rows, context_lens, scores, score_stride,
                                batch, num_q_heads, num_kv_heads, n_cols))
    return rc;
  if (int rc = check_select_args(who, n_cols, prev_budget, packed_slots, out_stride, sparse_lens)) return rc;
  SV_REQUIRE(tickets != nullptr, "%s: tickets scratch is null", who);
  SV_REQUIRE(page_table && page_table_stride >= n_cols + 1, "%s: page table (stride %lld) must cover n_cols+1=%d pages", who,
             (long long)page_table_stride, n_cols + 1);
  SV_REQUIRE(!refresh_bounds || page_rec, "%s: refresh_bounds needs page_rec", who);
  SV_REQUIRE(key && value && k_cache && v_cache && slot_mapping, "%s: null store pointer", who);
  SV_REQUIRE(aligned16(key) && aligned16(value) && aligned16(k_cache) && aligned16(v_cache),
             "%s: key/value/caches must be 16-byte aligned", who);
  SV_REQUIRE(k_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && k_stride % 8 == 0 &&
                 v_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && v_stride % 8 == 0,
             "%s: bad key/value token strides %lld/%lld", who, (long long)k_stride, (long long)v_stride);
  SV_REQUIRE(num_slots > 0 && (!refresh_bounds || num_slots % SVB200_PAGE == 0),
             "%s: num_slots=%lld must be positive (and a multiple of %d with refresh_bounds)", who, (long long)num_slots, SVB200_PAGE);
  const FusedStore st{key, k_stride, value, v_stride, k_cache, v_cache, slot_mapping, num_slots, refresh_bounds, page_rec};
  return launch_quest_score_select_pack(q, q_stride_b, bounds_layer, num_chunks, unit_table, unit_table_stride,
                                        page_table, page_table_stride,
                                        req_rows, context_lens, scores, score_stride, batch,
                                        num_q_heads, num_kv_heads, n_cols, prev_budget,
                                        packed_slots, out_stride, sparse_lens, selected_pages,
                                        tickets, &st, (cudaStream_t)stream);
}

static int check_comm(const char* who, const svb200_comm_t* comm) {
  SV_REQUIRE(comm->world >= 1 && comm->world <= SVB200_MAX_PEERS && comm->rank >= 0 && comm->rank < comm->world,
             "%s: comm world=%d rank=%d (world must be 1..%d)", who, comm->world, comm->rank, SVB200_MAX_PEERS);
  for (int r = 0; r < comm->world; ++r)
    SV_REQUIRE(comm->peer_buf[r] && comm->peer_sync[r] && aligned16(comm->peer_buf[r]),
               "%s: comm peer %d has a null / unaligned buffer", who, r);
  return 0;
}

static int fill_attn_params(AttnParams* p_out, const char* who, const void* q, int64_t q_stride_b, const void* k_cache,
                            const void* v_cache, int64_t num_slots, const int32_t* slot_table,
                            int64_t slot_table_stride, int32_t max_len, const int32_t* req_rows,
                            const int32_t* seq_lens, void* out, int64_t out_stride_b,
                            float* attn_score, int64_t score_stride, int32_t batch,
                            int32_t num_q_heads, int32_t num_kv_heads, int32_t heads_per_cta,
                            void* workspace, int64_t workspace_bytes, float sm_scale,
                            const svb200_comm_t* comm, int64_t gather_offset_bytes, int64_t gather_stride_b,
                            void* stream) {
  SV_REQUIRE(batch >= 0, "%s: batch=%d < 0", who, batch);
  if (batch == 0 && (comm == nullptr || comm->world <= 1)) return 0;
  SV_REQUIRE(batch > 0, "%s: an empty batch cannot take part in the tensor-parallel push", who);
  SV_REQUIRE(q && k_cache && v_cache && slot_table && req_rows && seq_lens && out, "%s: null pointer", who);
  SV_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && aligned16(out), "%s: q/k/v/out must be 16-byte aligned", who);
  if (int rc = check_heads(who, num_q_heads, num_kv_heads)) return rc;
  SV_REQUIRE(q_stride_b >= (int64_t)num_q_heads * SVB200_HEAD_DIM && q_stride_b % 8 == 0, "%s: bad q stride %lld", who, (long long)q_stride_b);
  SV_REQUIRE(out_stride_b >= (int64_t)num_q_heads * SVB200_HEAD_DIM && out_stride_b % 8 == 0, "%s: bad out stride %lld", who, (long long)out_stride_b);
  SV_REQUIRE(num_slots > 0 && slot_table_stride > 0, "%s: num_slots/slot_table_stride must be positive", who);
  SV_REQUIRE(max_len >= 0 && max_len <= slot_table_stride, "%s: max_len=%d must be in [0, slot_table_stride=%lld]", who, max_len, (long long)slot_table_stride);
  if (attn_score != nullptr)
    SV_REQUIRE(score_stride >= max_len, "%s: score_stride=%lld must cover max_len=%d when attn_score is given", who, (long long)score_stride, max_len);
  SV_REQUIRE(heads_per_cta == 0 || heads_per_cta == 1 || heads_per_cta == 2 || heads_per_cta == 4 || heads_per_cta == 8,
             "%s: heads_per_cta=%d must be 0 (auto), 1, 2, 4 or 8", who, heads_per_cta);
  const int64_t tk = ws_ticket_bytes(batch, num_kv_heads), per_cta = ws_cta_bytes(num_q_heads);
  SV_REQUIRE(workspace && aligned16(workspace) && workspace_bytes >= tk + per_cta,
             "%s: workspace must be 16-byte aligned and hold at least %lld bytes (svb200_workspace_bytes gives the full size)",
             who, (long long)(tk + per_cta));
  int64_t ctas = (workspace_bytes - tk) / per_cta;
  if (ctas > sm_count()) ctas = sm_count();
  AttnParams p;
  p.q = q; p.q_stride_b = q_stride_b; p.k_cache = k_cache; p.v_cache = v_cache;
  p.num_slots = num_slots; p.slot_table = slot_table; p.slot_table_stride = slot_table_stride;
  p.req_rows = req_rows; p.seq_lens = seq_lens; p.out = out;
  p.out_stride_b = out_stride_b; p.attn_score = attn_score; p.score_stride = score_stride;
  p.batch = batch; p.hq = num_q_heads; p.hkv = num_kv_heads; p.max_len = max_len;
  p.sm_scale = sm_scale; p.max_ctas = (int)ctas; p.stages = 0;
  char* w = static_cast<char*>(workspace);
  p.tickets = reinterpret_cast<int32_t*>(w);
  p.ws_lse = reinterpret_cast<float*>(w + tk);
  p.ws_o = reinterpret_cast<float*>(w + tk + ctas * 2 * num_q_heads * 4 * (int64_t)sizeof(float));
  p.no_math = 0;
  p.granule = 1;
  p.grid_cap = 0;
  p.evict_first = 1;
  p.push = PushArgs{};
  if (comm != nullptr && comm->world > 1) {
    if (int rc = check_comm(who, comm)) return rc;
    SV_REQUIRE(gather_offset_bytes >= 0 && gather_offset_bytes % 16 == 0 &&
                   gather_stride_b >= (int64_t)comm->world * num_q_heads * SVB200_HEAD_DIM && gather_stride_b % 8 == 0,
               "%s: gathered buffer offset %lld / row stride %lld (needs 16-byte alignment and >= world*Hq*128 elements per row)",
               who, (long long)gather_offset_bytes, (long long)gather_stride_b);
    p.push.world = comm->world;
    p.push.rank = comm->rank;
    p.push.stride_b = gather_stride_b;
    for (int r = 0; r < comm->world; ++r) {
      p.push.out[r] = reinterpret_cast<__nv_bfloat16*>(static_cast<char*>(comm->peer_buf[r]) + gather_offset_bytes);
      p.push.sync[r] = comm->peer_sync[r];
      p.push.ll[r] = static_cast<unsigned long long*>(comm->peer_ll[r]);
    }
    // flag-in-data staging, when the communicator has room for two areas of this launch's rows
    const int64_t words = (int64_t)comm->world * batch * num_q_heads * (SVB200_HEAD_DIM / 2);
    bool have_ll = comm->ll_bytes >= 2 * words * 8;
    for (int r = 0; r < comm->world && have_ll; ++r) have_ll = comm->peer_ll[r] != nullptr && aligned16(comm->peer_ll[r]);
    p.push.ll_words = have_ll ? words : 0;
  }
  *p_out = p;
  return 0;
}

static int decode_attn_impl(const char* who, const void* q, int64_t q_stride_b, const void* k_cache,
                            const void* v_cache, int64_t num_slots, const int32_t* slot_table,
                            int64_t slot_table_stride, int32_t max_len, const int32_t* req_rows,
                            const int32_t* seq_lens, void* out, int64_t out_stride_b,
                            float* attn_score, int64_t score_stride, int32_t batch,
                            int32_t num_q_heads, int32_t num_kv_heads, int32_t heads_per_cta,
                            void* workspace, int64_t workspace_bytes, float sm_scale,
                            const svb200_comm_t* comm, int64_t gather_offset_bytes, int64_t gather_stride_b,
                            void* stream) {
  if (batch == 0 && (comm == nullptr || comm->world <= 1)) return 0;
  AttnParams p;
  if (int rc = fill_attn_params(&p, who, q, q_stride_b, k_cache, v_cache, num_slots, slot_table, slot_table_stride,
                                max_len, req_rows, seq_lens, out, out_stride_b, attn_score, score_stride, batch,
                                num_q_heads, num_kv_heads, heads_per_cta, workspace, workspace_bytes, sm_scale, comm,
                                gather_offset_bytes, gather_stride_b, stream))
    return rc;
  return launch_sparse_decode_attn(p, heads_per_cta, (cudaStream_t)stream);
}

int svb200_quest_decode_layer(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                              const int32_t* slot_mapping, int32_t refresh_bounds, const int32_t* page_rec,
                              void* k_cache, void* v_cache, int64_t num_slots,
                              const void* q, int64_t q_stride_b, void* bounds_layer,
                              int64_t num_chunks, const int32_t* unit_table, int64_t unit_table_stride,
                              const int32_t* page_table, int64_t page_table_stride,
                              const int32_t* req_rows, const int32_t* context_lens,
                              void* scores, int64_t score_stride, int32_t batch,
                              int32_t num_q_heads, int32_t num_kv_heads, int32_t n_cols,
                              int32_t prev_budget, int32_t* packed_slots, int64_t out_stride,
                              int32_t* sparse_lens, int32_t* selected_pages, int32_t* sched_scratch,
                              void* out, int64_t out_stride_b, void* workspace, int64_t workspace_bytes,
                              float sm_scale, const svb200_comm_t* comm, int64_t gather_offset_bytes,
                              int64_t gather_stride_b, void* stream) {
  const char* who = "quest_decode_layer";
  SV_REQUIRE(batch > 0, "%s: batch=%d must be positive", who, batch);
  if (int rc = check_score_args(who, q, q_stride_b, bounds_layer, num_chunks, unit_table,
                                unit_table_stride, req_rows, context_lens, scores, score_stride,
                                batch, num_q_heads, num_kv_heads, n_cols))
    return rc;
  if (int rc = check_select_args(who, n_cols, prev_budget, packed_slots, out_stride, sparse_lens)) return rc;
  SV_REQUIRE(sched_scratch != nullptr, "%s: scheduler scratch is null", who);
  SV_REQUIRE(page_table && page_table_stride >= n_cols + 1, "%s: page table (stride %lld) must cover n_cols+1=%d pages", who,
             (long long)page_table_stride, n_cols + 1);
  SV_SUPPORTED(num_q_heads / num_kv_heads <= 8 && n_cols > 0,
                 "%s: needs n_cols > 0 and at most 8 q heads per KV head; use the separate entry points", who);
  FusedStore st{};
  if (key != nullptr) {
    SV_REQUIRE(!refresh_bounds || page_rec, "%s: refresh_bounds needs page_rec", who);
    SV_REQUIRE(value && slot_mapping, "%s: null store pointer", who);
    SV_REQUIRE(aligned16(key) && aligned16(value), "%s: key/value must be 16-byte aligned", who);
    SV_REQUIRE(k_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && k_stride % 8 == 0 &&
                   v_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && v_stride % 8 == 0,
               "%s: bad key/value token strides %lld/%lld", who, (long long)k_stride, (long long)v_stride);
    SV_REQUIRE(num_slots > 0 && (!refresh_bounds || num_slots % SVB200_PAGE == 0),
               "%s: num_slots=%lld must be positive (and a multiple of %d with refresh_bounds)", who, (long long)num_slots, SVB200_PAGE);
    st = FusedStore{key, k_stride, value, v_stride, k_cache, v_cache, slot_mapping, num_slots, refresh_bounds, page_rec};
  }
  AttnParams ap;
  // the attention phase reads the slots this launch packs: row b of packed_slots, (prev_budget + 1) pages at most
  if (int rc = fill_attn_params(&ap, who, q, q_stride_b, k_cache, v_cache, num_slots, packed_slots, out_stride,
                                (prev_budget + 1) * SVB200_PAGE, context_lens /* placeholders: the phase takes the */,
                                context_lens /* identity rows and derives the lengths from context_lens */, out,
                                out_stride_b, nullptr, 0, batch, num_q_heads, num_kv_heads, 0, workspace,
                                workspace_bytes, sm_scale, comm, gather_offset_bytes, gather_stride_b, stream))
    return rc;
  ap.req_rows = nullptr;
  ap.seq_lens = nullptr;
  return launch_quest_decode_layer(q, q_stride_b, bounds_layer, num_chunks, unit_table, unit_table_stride, page_table,
                                   page_table_stride, req_rows, context_lens, scores, score_stride, batch, num_q_heads,
                                   num_kv_heads, n_cols, prev_budget, packed_slots, out_stride, sparse_lens,
                                   selected_pages, sched_scratch, key != nullptr ? &st : nullptr, ap,
                                   (cudaStream_t)stream);
}

int svb200_sparse_decode_attn(const void* q, int64_t q_stride_b, const void* k_cache,
                              const void* v_cache, int64_t num_slots, const int32_t* slot_table,
                              int64_t slot_table_stride, int32_t max_len, const int32_t* req_rows,
                              const int32_t* seq_lens, void* out, int64_t out_stride_b,
                              float* attn_score, int64_t score_stride, int32_t batch,
                              int32_t num_q_heads, int32_t num_kv_heads, int32_t heads_per_cta,
                              void* workspace, int64_t workspace_bytes, float sm_scale,
                              void* stream) {
  return decode_attn_impl("sparse_decode_attn", q, q_stride_b, k_cache, v_cache, num_slots, slot_table,
                          slot_table_stride, max_len, req_rows, seq_lens, out, out_stride_b, attn_score,
                          score_stride, batch, num_q_heads, num_kv_heads, heads_per_cta, workspace,
                          workspace_bytes, sm_scale, nullptr, 0, 0, stream);
}

int svb200_sparse_decode_attn_tp(const void* q, int64_t q_stride_b, const void* k_cache,
                                 const void* v_cache, int64_t num_slots, const int32_t* slot_table,
                                 int64_t slot_table_stride, int32_t max_len, const int32_t* req_rows,
                                 const int32_t* seq_lens, void* out, int64_t out_stride_b,
                                 float* attn_score, int64_t score_stride, int32_t batch,
                                 int32_t num_q_heads, int32_t num_kv_heads, int32_t heads_per_cta,
                                 void* workspace, int64_t workspace_bytes, float sm_scale,
                                 const svb200_comm_t* comm, int64_t gather_offset_bytes,
                                 int64_t gather_stride_b, void* stream) {
  const char* who = "sparse_decode_attn_tp";
  SV_REQUIRE(comm != nullptr, "%s: comm is null", who);
  return decode_attn_impl(who, q, q_stride_b, k_cache, v_cache, num_slots, slot_table,
                          slot_table_stride, max_len, req_rows, seq_lens, out, out_stride_b, attn_score,
                          score_stride, batch, num_q_heads, num_kv_heads, heads_per_cta, workspace,
                          workspace_bytes, sm_scale, comm, gather_offset_bytes, gather_stride_b, stream);
}

int svb200_allreduce_sum_bf16(const svb200_comm_t* comm, int64_t in_offset_bytes, int64_t n_elems,
                              void* out, void* stream) {
  const char* who = "allreduce_sum_bf16";
  SV_REQUIRE(comm != nullptr && out != nullptr, "%s: null pointer", who);
  if (int rc = check_comm(who, comm)) return rc;
  SV_REQUIRE(n_elems >= 0 && n_elems % 8 == 0 && in_offset_bytes >= 0 && in_offset_bytes % 16 == 0 && aligned16(out),
             "%s: n_elems=%lld must be a multiple of 8, offsets / out 16-byte aligned", who, (long long)n_elems);
  if (n_elems == 0) return 0;
  AllReduceArgs a{};
  a.world = comm->world;
  a.rank = comm->rank;
  for (int r = 0; r < comm->world; ++r) {
    a.in[r] = reinterpret_cast<const __nv_bfloat16*>(static_cast<const char*>(comm->peer_buf[r]) + in_offset_bytes);
    a.sync[r] = comm->peer_sync[r];
  }
  a.out = static_cast<__nv_bfloat16*>(out);
  a.n = n_elems;
  return launch_allreduce_sum_bf16(a, (cudaStream_t)stream);
}

/* ---- peer-mapped ("symmetric") memory: plain cudaMalloc + CUDA IPC handles ---- */
int svb200_symm_alloc(int64_t bytes, void** ptr_host, void* handle64_host) {
  SV_REQUIRE(bytes > 0 && ptr_host && handle64_host, "symm_alloc: bad arguments");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, (size_t)bytes);
  if (e != cudaSuccess) return cuda_error("symm_alloc cudaMalloc", e);
  e = cudaMemset(p, 0, (size_t)bytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(static_cast<cudaIpcMemHandle_t*>(handle64_host), p);
  if (e != cudaSuccess) {
    cudaFree(p);
    return cuda_error("symm_alloc", e);
  }
  *ptr_host = p;
  return 0;
}
int svb200_symm_free(void* ptr) {
  const cudaError_t e = cudaFree(ptr);
  return e == cudaSuccess ? 0 : cuda_error("symm_free", e);
}
int svb200_ipc_open(const void* handle64_host, void** ptr_host) {
  SV_REQUIRE(handle64_host && ptr_host, "ipc_open: null pointer");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle64_host, sizeof(h));
  void* p = nullptr;
  const cudaError_t e = cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess);
  if (e != cudaSuccess) return cuda_error("ipc_open", e);
  *ptr_host = p;
  return 0;
}
int svb200_ipc_close(void* ptr) {
  const cudaError_t e = cudaIpcCloseMemHandle(ptr);
  return e == cudaSuccess ? 0 : cuda_error("ipc_close", e);
}

int svb200_h2o_normalize_accumulate(float* score, int64_t score_stride, int32_t batch,
                                    int32_t width, float scale, float* cumulative,
                                    int64_t cum_stride, const int32_t* cum_rows, const int32_t* kv_lens,
                                    float refill, void* stream) {
  SV_REQUIRE(batch >= 0 && width >= 0, "h2o_normalize_accumulate: negative batch/width");
  if (batch == 0 || width == 0) return 0;
  SV_REQUIRE(score != nullptr && score_stride >= width, "h2o_normalize_accumulate: score null or stride %lld < width %d", (long long)score_stride, width);
  if (cumulative != nullptr)
    SV_REQUIRE(kv_lens != nullptr && cum_stride >= 1, "h2o_normalize_accumulate: cumulative needs kv_lens and a positive stride");
  return launch_h2o_normalize_accumulate(score, score_stride, batch, width, scale, cumulative,
                                         cum_stride, cum_rows, kv_lens, refill, (cudaStream_t)stream);
}

int svb200_h2o_prepare_decode(const int32_t* meta, const int32_t* free_stack, int64_t free_stride,
                              int32_t* slot_tables, int64_t layer_stride, int64_t row_stride,
                              int32_t* layer_slot_mapping, int64_t lsm_stride, int32_t* context_lens,
                              int32_t* req_indices, int32_t* slot_mapping, int32_t* kv_lens, int32_t* rows_out,
                              int32_t batch, int32_t num_layers, void* stream) {
  SV_REQUIRE(batch >= 0 && num_layers >= 0, "h2o_prepare_decode: negative batch / layers");
  if (batch == 0 || num_layers == 0) return 0;
  SV_REQUIRE(meta && free_stack && slot_tables && layer_slot_mapping && context_lens && req_indices && slot_mapping &&
                 kv_lens && rows_out, "h2o_prepare_decode: null pointer");
  SV_REQUIRE(free_stride > 0 && layer_stride > 0 && row_stride > 0 && lsm_stride >= batch, "h2o_prepare_decode: bad strides");
  return launch_h2o_prepare_decode(meta, free_stack, free_stride, slot_tables, layer_stride, row_stride, layer_slot_mapping,
                                   lsm_stride, context_lens, req_indices, slot_mapping, kv_lens, rows_out, batch, num_layers,
                                   (cudaStream_t)stream);
}

int svb200_prefill_attn(const void* q, int64_t q_stride, const void* k_cache, const void* v_cache, int64_t num_slots,
                        const int32_t* slot_table, int64_t slot_table_stride, const int32_t* req_rows,
                        const int32_t* start_loc, const int32_t* context_lens, const int32_t* prompt_cache_lens,
                        void* out, int64_t out_stride, int32_t batch, int32_t num_q_heads, int32_t num_kv_heads,
                        int32_t max_chunk_len, float sm_scale, void* stream) {
  const char* who = "prefill_attn";
  SV_REQUIRE(batch >= 0 && max_chunk_len >= 0, "%s: negative sizes", who);
  if (batch == 0 || max_chunk_len == 0) return 0;
  SV_REQUIRE(q && k_cache && v_cache && slot_table && req_rows && start_loc && context_lens && prompt_cache_lens && out,
             "%s: null pointer", who);
  SV_REQUIRE(aligned16(q) && aligned16(k_cache) && aligned16(v_cache) && aligned16(out), "%s: q/k/v/out must be 16-byte aligned", who);
  if (int rc = check_heads(who, num_q_heads, num_kv_heads)) return rc;
  SV_SUPPORTED(num_q_heads / num_kv_heads <= 8, "%s: GQA group size %d > 8 is not supported", who, num_q_heads / num_kv_heads);
  SV_REQUIRE(q_stride >= (int64_t)num_q_heads * SVB200_HEAD_DIM && q_stride % 8 == 0 &&
                 out_stride >= (int64_t)num_q_heads * SVB200_HEAD_DIM && out_stride % 2 == 0,
             "%s: bad q / out token strides %lld / %lld", who, (long long)q_stride, (long long)out_stride);
  SV_REQUIRE(num_slots > 0 && slot_table_stride > 0, "%s: num_slots / slot_table_stride must be positive", who);
  SV_REQUIRE(batch * (int64_t)num_kv_heads <= 65535, "%s: batch * kv heads = %lld exceeds the grid's y extent", who,
             (long long)batch * num_kv_heads);
  PrefillAttnHostArgs h{q, q_stride, k_cache, v_cache, num_slots, slot_table, slot_table_stride, req_rows, start_loc,
                        context_lens, prompt_cache_lens, out, out_stride, batch, num_q_heads, num_kv_heads, max_chunk_len,
                        sm_scale};
  return launch_prefill_attn(h, (cudaStream_t)stream);
}

int svb200_h2o_prefill_score(const void* q, int64_t q_stride, const void* k_cache, int64_t num_slots,
                             const int32_t* slot_table, int64_t slot_table_stride, const int32_t* req_rows,
                             const int32_t* start_loc, const int32_t* context_lens,
                             const int32_t* prompt_cache_lens, const int32_t* score_starts,
                             const int32_t* score_ends, float* score, int64_t score_stride,
                             float* cumulative, int64_t cum_stride, const int32_t* cum_rows, int32_t batch,
                             int32_t num_q_heads, int32_t num_kv_heads, int32_t max_query_len,
                             int32_t max_context_len, int32_t candidate_start, int32_t num_recent_tokens,
                             float sm_scale, void* stream) {
  const char* who = "h2o_prefill_score";
  SV_REQUIRE(batch >= 0 && max_query_len >= 0 && max_context_len >= 0, "%s: negative sizes", who);
  if (batch == 0 || max_query_len == 0 || max_context_len == 0) return 0;
  SV_REQUIRE(q && k_cache && slot_table && req_rows && start_loc && context_lens && prompt_cache_lens && score_starts &&
                 score_ends && score, "%s: null pointer", who);
  SV_REQUIRE(aligned16(q) && aligned16(k_cache), "%s: q / k_cache must be 16-byte aligned", who);
  if (int rc = check_heads(who, num_q_heads, num_kv_heads)) return rc;
  SV_REQUIRE(q_stride >= (int64_t)num_q_heads * SVB200_HEAD_DIM && q_stride % 8 == 0, "%s: bad q token stride %lld", who, (long long)q_stride);
  SV_SUPPORTED(max_query_len <= 128, "%s: probability scoring serves windows of <= 128 queries, got %d (prefill_score.py:505-509)", who, max_query_len);
  SV_REQUIRE(score_stride >= max_context_len && slot_table_stride >= max_context_len && num_slots > 0,
             "%s: score_stride=%lld / slot_table_stride=%lld must cover max_context_len=%d", who, (long long)score_stride,
             (long long)slot_table_stride, max_context_len);
  SV_REQUIRE(candidate_start >= 0 && num_recent_tokens >= 0, "%s: negative candidate range", who);
  if (cumulative != nullptr) SV_REQUIRE(cum_rows != nullptr && cum_stride >= max_context_len, "%s: cumulative needs cum_rows and a stride covering max_context_len", who);
  PrefillScoreHostArgs h{q, q_stride, k_cache, num_slots, slot_table, slot_table_stride, req_rows, start_loc, context_lens,
                         prompt_cache_lens, score_starts, score_ends, score, score_stride, cumulative, cum_stride, cum_rows,
                         batch, num_q_heads, num_kv_heads, max_query_len, max_context_len, candidate_start, num_recent_tokens,
                         sm_scale};
  return launch_h2o_prefill_score(h, (cudaStream_t)stream);
}

int svb200_h2o_select_compact(float* cumulative, const int64_t* cum_rows, int32_t* slot_tables,
                              const int64_t* slot_rows, int32_t* free_stack,
                              const int64_t* free_rows, int32_t* keep_out, int32_t num_rows,
                              const int32_t* kv_lens, int32_t kv_len, int32_t budget, int32_t recent,
                              void* stream) {
  SV_REQUIRE(num_rows >= 0, "h2o_select_compact: num_rows=%d < 0", num_rows);
  if (num_rows == 0) return 0;
  SV_REQUIRE(cumulative && cum_rows && slot_tables && slot_rows, "h2o_select_compact: null pointer");
  SV_REQUIRE(budget > 0 && kv_len > budget, "h2o_select_compact: need kv_len=%d > budget=%d > 0", kv_len, budget);
  SV_REQUIRE(recent >= 1 && recent <= budget, "h2o_select_compact: recent=%d must be in [1, budget=%d]", recent, budget);
  SV_REQUIRE((free_stack == nullptr) == (free_rows == nullptr), "h2o_select_compact: free_stack and free_rows go together");
  SV_SUPPORTED((int64_t)kv_len * 12 <= 200 * 1024, "h2o_select_compact: kv_len=%d too long for the shared-memory row cache", kv_len);
  return launch_h2o_select_compact(cumulative, cum_rows, slot_tables, slot_rows, free_stack,
                                   free_rows, keep_out, num_rows, kv_lens, kv_len, budget, recent,
                                   (cudaStream_t)stream);
}

int svb200_store_kv_minmax(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                           void* k_cache, void* v_cache, const int32_t* slot_mapping,
                           int32_t n_tokens, int32_t num_kv_heads, int64_t num_slots,
                           void* quest_bounds_layer, const int32_t* page_rec, int64_t num_chunks,
                           void* stream) {
  SV_REQUIRE(n_tokens >= 0, "store_kv_minmax: n_tokens=%d < 0", n_tokens);
  if (n_tokens == 0) return 0;
  SV_REQUIRE(key && value && k_cache && v_cache && slot_mapping, "store_kv_minmax: null pointer");
  SV_REQUIRE(aligned16(key) && aligned16(value) && aligned16(k_cache) && aligned16(v_cache),
             "store_kv_minmax: buffers must be 16-byte aligned");
  SV_REQUIRE(num_kv_heads > 0 && k_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && k_stride % 8 == 0 &&
                 v_stride >= (int64_t)num_kv_heads * SVB200_HEAD_DIM && v_stride % 8 == 0,
             "store_kv_minmax: bad key/value token strides %lld/%lld", (long long)k_stride, (long long)v_stride);
  SV_REQUIRE(num_slots > 0, "store_kv_minmax: num_slots must be positive");
  if (quest_bounds_layer != nullptr)
    SV_REQUIRE(aligned16(quest_bounds_layer) && num_slots % SVB200_PAGE == 0 && page_rec && num_chunks > 0,
               "store_kv_minmax: bounds need 16-byte alignment, paged slots, page_rec and num_chunks");
  return launch_store_kv_minmax(key, k_stride, value, v_stride, k_cache, v_cache, slot_mapping,
                                n_tokens, num_kv_heads, num_slots, quest_bounds_layer, page_rec, num_chunks,
                                (cudaStream_t)stream);
}

}  // extern "C"
