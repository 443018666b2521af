// One decode token's K / V rows into the paged cache, with the optional fused Quest bounds refresh
// (shared by the stand-alone store kernel and the store fused into the page-scoring launch).
//
// Reference behaviour: kernels/triton/store_kvcache.py:10-71 (scatter one K and one V row per token,
// slot -1 skipped) and engine/cache_manager/quest.py:1719-1771 (bounds of a page are rebuilt when a
// decode token completes it).
#pragma once

#include "bounds_layout.cuh"
#include "common.cuh"

namespace svb200 {

struct StoreArgs {
  const __nv_bfloat16* key;  // [tokens, Hkv, 128], token stride k_stride elements
  int64_t k_stride;
  const __nv_bfloat16* value;
  int64_t v_stride;
  __nv_bfloat16* k_cache;  // [num_slots, Hkv, 128]
  __nv_bfloat16* v_cache;
  const int32_t* slot_mapping;  // [tokens]
  int64_t num_slots;
  __nv_bfloat16* bounds;  // bounds chunks of this layer (bounds_layout.cuh), or null: no refresh
  const int32_t* page_rec;  // physical page -> bounds record (-1: the page belongs to no row), with bounds
  int64_t num_recs;         // 16 * number of bounds chunks
};

// Thread `t` of `nthr` moves every nthr-th 16-byte vector of token `tok`.  Bounds refresh: valid when no
// two tokens of the launch share a page (decode: one token per sequence): rows 0..14 of the page are
// then already in the cache and row 15 is the vector this thread holds.
__device__ __forceinline__ void store_kv_token(const StoreArgs& a, int tok, int hkv, int t, int nthr) {
  const int slot = a.slot_mapping[tok];
  if (slot < 0 || slot >= a.num_slots) return;
  const int64_t row_elems = (int64_t)hkv * kHeadDim;
  const int vecs = hkv * (kHeadDim / 8);
  for (int v = t; v < vecs; v += nthr) {
    const uint4 kv = *reinterpret_cast<const uint4*>(a.key + tok * a.k_stride + v * 8);
    const uint4 vv = *reinterpret_cast<const uint4*>(a.value + tok * a.v_stride + v * 8);
    *reinterpret_cast<uint4*>(a.k_cache + slot * row_elems + v * 8) = kv;
    *reinterpret_cast<uint4*>(a.v_cache + slot * row_elems + v * 8) = vv;
    if (a.bounds != nullptr && (slot & (kPage - 1)) == kPage - 1) {
      const int64_t page = slot >> 4;
      const __nv_bfloat16* page_base = a.k_cache + page * kPage * row_elems + v * 8;
      uint4 mx = kv, mn = kv;
      __nv_bfloat162* mxp = reinterpret_cast<__nv_bfloat162*>(&mx);
      __nv_bfloat162* mnp = reinterpret_cast<__nv_bfloat162*>(&mn);
      // five rows in flight at a time (fifteen would cost 60 registers in the scoring kernel)
#pragma unroll 1
      for (int r0 = 0; r0 < kPage - 1; r0 += 5) {
        uint4 rows[5];
#pragma unroll
        for (int r = 0; r < 5; ++r) rows[r] = *reinterpret_cast<const uint4*>(page_base + (r0 + r) * row_elems);
#pragma unroll
        for (int r = 0; r < 5; ++r) {
          const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&rows[r]);
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            mxp[i] = __hmax2(mxp[i], x[i]);
            mnp[i] = __hmin2(mnp[i], x[i]);
          }
        }
      }
      const int rec = a.page_rec[page];
      if (rec >= 0 && rec < a.num_recs) {
        const int hs = bounds_head_split(hkv);
        *reinterpret_cast<uint4*>(a.bounds + bounds_seg_offset(rec, 0, v >> 4, v & 15, hkv, hs)) = mx;
        *reinterpret_cast<uint4*>(a.bounds + bounds_seg_offset(rec, 1, v >> 4, v & 15, hkv, hs)) = mn;
      }
    }
  }
}

}  // namespace svb200
