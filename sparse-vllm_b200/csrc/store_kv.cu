// KV-row store for sm_100a with an optional fused Quest page-bounds refresh.
//
// Reference behaviour (CURRENTF/Sparse-vLLM): kernels/triton/store_kvcache.py:10-71
// (scatter one K and one V row per token, slot -1 skipped) and, for the fused
// refresh, engine/cache_manager/quest.py:1719-1771 (bounds of a page are rebuilt
// when a decode token completes it).
#include "common.cuh"
#include "launch.h"

namespace svb200 {

// One CTA per token; thread v moves the v-th 16-byte vector of the K and V rows.
// Fused bounds refresh: valid when no two tokens of the launch share a page
// (decode: one token per sequence) — rows 0..14 of the page are then already in
// the cache and row 15 is the vector this thread just wrote.
__global__ void __launch_bounds__(128) store_kv_minmax_kernel(
    const __nv_bfloat16* __restrict__ key, int64_t k_stride, const __nv_bfloat16* __restrict__ value,
    int64_t v_stride, __nv_bfloat16* __restrict__ k_cache, __nv_bfloat16* __restrict__ v_cache,
    const int32_t* __restrict__ slot_mapping, int hkv, int64_t num_slots,
    __nv_bfloat16* __restrict__ bounds) {
  const int tok = blockIdx.x;
  const int slot = slot_mapping[tok];
  if (slot < 0 || slot >= num_slots) return;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const int vecs = hkv * (kHeadDim / 8);
  for (int v = threadIdx.x; v < vecs; v += blockDim.x) {
    const uint4 kv = *reinterpret_cast<const uint4*>(key + tok * k_stride + v * 8);
    const uint4 vv = *reinterpret_cast<const uint4*>(value + tok * v_stride + v * 8);
    *reinterpret_cast<uint4*>(k_cache + slot * row_elems + v * 8) = kv;
    *reinterpret_cast<uint4*>(v_cache + slot * row_elems + v * 8) = vv;
    if (bounds != nullptr && (slot & (kPage - 1)) == kPage - 1) {
      const int64_t page = slot >> 4;
      const __nv_bfloat16* page_base = k_cache + page * kPage * row_elems + v * 8;
      uint4 mx = kv, mn = kv;
      __nv_bfloat162* mxp = reinterpret_cast<__nv_bfloat162*>(&mx);
      __nv_bfloat162* mnp = reinterpret_cast<__nv_bfloat162*>(&mn);
      uint4 rows[kPage - 1];
#pragma unroll
      for (int r = 0; r < kPage - 1; ++r)
        rows[r] = *reinterpret_cast<const uint4*>(page_base + r * row_elems);
#pragma unroll
      for (int r = 0; r < kPage - 1; ++r) {
        const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&rows[r]);
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          mxp[i] = __hmax2(mxp[i], x[i]);
          mnp[i] = __hmin2(mnp[i], x[i]);
        }
      }
      __nv_bfloat16* out = bounds + page * 2 * row_elems;
      *reinterpret_cast<uint4*>(out + v * 8) = mx;
      *reinterpret_cast<uint4*>(out + row_elems + v * 8) = mn;
    }
  }
}

int launch_store_kv_minmax(const void* key, int64_t k_stride, const void* value, int64_t v_stride,
                           void* k_cache, void* v_cache, const int32_t* slot_mapping, int n_tokens,
                           int hkv, int64_t num_slots, void* bounds_layer, cudaStream_t stream) {
  if (n_tokens <= 0) return 0;
  store_kv_minmax_kernel<<<n_tokens, 128, 0, stream>>>(
      (const __nv_bfloat16*)key, k_stride, (const __nv_bfloat16*)value, v_stride,
      (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, slot_mapping, hkv, num_slots,
      (__nv_bfloat16*)bounds_layer);
  return check_launch("store_kv_minmax");
}

}  // namespace svb200
