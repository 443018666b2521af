// KV-row store for sm_100a with an optional fused Quest page-bounds refresh.
//
// Reference behaviour (CURRENTF/Sparse-vLLM): kernels/triton/store_kvcache.py:10-71
// (scatter one K and one V row per token, slot -1 skipped) and, for the fused
// refresh, engine/cache_manager/quest.py:1719-1771 (bounds of a page are rebuilt
// when a decode token completes it).
#include "common.cuh"
#include "launch.h"
#include "store_kv.cuh"

namespace svb200 {

// One CTA per token (store_kv.cuh).
__global__ void __launch_bounds__(128) store_kv_minmax_kernel(StoreArgs a, int hkv) {
  store_kv_token(a, blockIdx.x, hkv, threadIdx.x, blockDim.x);
}

int launch_store_kv_minmax(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                           void* k_cache, void* v_cache, const int32_t* slot_mapping, int n_tokens,
                           int hkv, int64_t num_slots, void* bounds_layer, const int32_t* page_rec,
                           int64_t num_chunks, cudaStream_t stream) {
  if (n_tokens <= 0) return 0;
  const StoreArgs a{(const __nv_bfloat16*)key, k_stride, (const __nv_bfloat16*)value, v_stride,
                    (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, slot_mapping, num_slots,
                    (__nv_bfloat16*)bounds_layer, page_rec, num_chunks * kPage};
  store_kv_minmax_kernel<<<n_tokens, 128, 0, stream>>>(a, hkv);
  return check_launch("store_kv_minmax");
}

}  // namespace svb200
