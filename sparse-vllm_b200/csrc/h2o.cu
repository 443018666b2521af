// H2O kernels for sm_100a: layer-end score normalisation with fused cumulative
// accumulation, and burst eviction (heavy-hitter selection + index-only compaction).
//
// Reference behaviour (CURRENTF/Sparse-vLLM, src/sparsevllm/engine/):
//   sparse_controller.py:761-767         score *= D^-0.5 ; softmax over the buffer width
//   cache_manager/h2o.py:958-1038        cumulative[:kv_len] = step + pad(previous, 1)
//   cache_manager/h2o.py:519-563         heavy hitters (stable descending) + recent suffix
//   cache_manager/snapkv.py:1681-1803    slot-row compaction + free-stack push
#include "common.cuh"
#include "launch.h"

namespace svb200 {

constexpr int kH2oThreads = 256;

__device__ __forceinline__ float block_reduce_max(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = lane < nwarps ? red[lane] : -INFINITY;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r = fmaxf(r, __shfl_xor_sync(0xffffffffu, r, o));
  return r;
}
__device__ __forceinline__ float block_reduce_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if (lane == 0) red[warp] = v;
  __syncthreads();
  float r = lane < nwarps ? red[lane] : 0.f;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) r += __shfl_xor_sync(0xffffffffu, r, o);
  return r;
}

// One CTA per row.  The scaled row is cached in shared memory when it fits.
template <bool CACHE_ROW>
__global__ void __launch_bounds__(kH2oThreads) h2o_normalize_accumulate_kernel(
    float* __restrict__ score, int64_t score_stride, int width, float scale,
    float* __restrict__ cumulative, int64_t cum_stride, const int32_t* __restrict__ kv_lens) {
  extern __shared__ float row_cache[];
  __shared__ float red[32];
  const int b = blockIdx.x, tid = threadIdx.x;
  float* row = score + (int64_t)b * score_stride;

  float mx = -INFINITY;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float x = row[i] * scale;
    if (CACHE_ROW) row_cache[i] = x;
    mx = fmaxf(mx, x);
  }
  mx = block_reduce_max(mx, red);
  float sum = 0.f;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float x = CACHE_ROW ? row_cache[i] : row[i] * scale;
    const float e = expf(x - mx);
    if (CACHE_ROW) row_cache[i] = e;
    sum += e;
  }
  sum = block_reduce_sum(sum, red);
  const int kv_len = (cumulative != nullptr && kv_lens != nullptr) ? min(kv_lens[b], width) : 0;
  float* cum = cumulative != nullptr ? cumulative + (int64_t)b * cum_stride : nullptr;
  for (int i = tid; i < width; i += kH2oThreads) {
    const float e = CACHE_ROW ? row_cache[i] : expf(row[i] * scale - mx);
    const float pr = e / sum;
    row[i] = pr;
    if (i < kv_len) cum[i] = (i < kv_len - 1) ? cum[i] + pr : pr;
  }
}

// fp32 bits -> 32-bit key that orders like the float value.
__device__ __forceinline__ uint32_t f32_order_key(uint32_t bits) {
  return (bits & 0x80000000u) ? ~bits : (bits | 0x80000000u);
}

__device__ __forceinline__ int block_exclusive_scan_i(int v, int* warp_scan) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
  int inc = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int n = __shfl_up_sync(0xffffffffu, inc, o);
    if (lane >= o) inc += n;
  }
  __syncthreads();
  if (lane == 31) warp_scan[warp] = inc;
  __syncthreads();
  if (warp == 0) {
    const int w = lane < nwarps ? warp_scan[lane] : 0;
    int winc = w;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, winc, o);
      if (lane >= o) winc += n;
    }
    warp_scan[lane] = winc - w;
  }
  __syncthreads();
  return warp_scan[warp] + inc - v;
}

// One CTA per (layer,row) entry.  smem: keys[kv_len] u32, vals[kv_len] f32, slots[kv_len] i32.
__global__ void __launch_bounds__(kH2oThreads) h2o_select_compact_kernel(
    float* __restrict__ cumulative, const int64_t* __restrict__ cum_rows,
    int32_t* __restrict__ slot_tables, const int64_t* __restrict__ slot_rows,
    int32_t* __restrict__ free_stack, const int64_t* __restrict__ free_rows,
    int32_t* __restrict__ keep_out, int kv_len, int budget, int recent) {
  extern __shared__ __align__(16) unsigned char h2o_smem[];
  uint32_t* keys = reinterpret_cast<uint32_t*>(h2o_smem);
  float* vals = reinterpret_cast<float*>(keys + kv_len);
  int32_t* slots = reinterpret_cast<int32_t*>(vals + kv_len);
  __shared__ int hist[256];
  __shared__ int warp_scan[32];
  __shared__ int bcast[2];

  const int r = blockIdx.x, tid = threadIdx.x;
  float* cum = cumulative + cum_rows[r];
  int32_t* srow = slot_tables + slot_rows[r];
  const int recent_start = kv_len - recent;
  const int heavy = budget - recent;

  for (int i = tid; i < kv_len; i += kH2oThreads) {
    const float v = cum[i];
    vals[i] = v;
    keys[i] = f32_order_key(__float_as_uint(v));
    slots[i] = srow[i];
  }
  __syncthreads();

  // ---- radix select of the heavy-th largest key among [0, recent_start)
  uint32_t prefix = 0, prefix_mask = 0;
  int k = heavy;
  if (heavy > 0) {
    for (int shift = 24; shift >= 0; shift -= 8) {
      for (int i = tid; i < 256; i += kH2oThreads) hist[i] = 0;
      __syncthreads();
      for (int i = tid; i < recent_start; i += kH2oThreads) {
        const uint32_t key = keys[i];
        if ((key & prefix_mask) == prefix) atomicAdd(&hist[(key >> shift) & 0xff], 1);
      }
      __syncthreads();
      if (tid < 32) {
        const int lane = tid;
        int local[8], lsum = 0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          local[i] = hist[lane * 8 + i];
          lsum += local[i];
        }
        int suf = lsum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const int n = __shfl_down_sync(0xffffffffu, suf, o);
          if (lane + o < 32) suf += n;
        }
        const int above_lane = suf - lsum;
        if (above_lane < k && k <= suf) {
          int above = above_lane, bin = lane * 8 + 7;
#pragma unroll
          for (int i = 7; i >= 0; --i) {
            if (above + local[i] >= k) {
              bin = lane * 8 + i;
              break;
            }
            above += local[i];
          }
          bcast[0] = bin;
          bcast[1] = above;
        }
      }
      __syncthreads();
      prefix |= (uint32_t)bcast[0] << shift;
      prefix_mask |= 0xffu << shift;
      k -= bcast[1];
      __syncthreads();
    }
  }
  const uint32_t thresh = prefix;  // key of the heavy-th largest; k = ties to take (lowest index first)

  // ---- ordered compaction over contiguous chunks of [0, kv_len)
  const int per = (kv_len + kH2oThreads - 1) / kH2oThreads;
  const int i0 = min(tid * per, kv_len), i1 = min(i0 + per, kv_len);
  int cnt_eq = 0, cnt_sure = 0;
  for (int i = i0; i < i1; ++i) {
    if (i >= recent_start) {
      ++cnt_sure;
    } else if (heavy > 0) {
      cnt_sure += keys[i] > thresh;
      cnt_eq += keys[i] == thresh;
    }
  }
  const int eq_before = block_exclusive_scan_i(cnt_eq, warp_scan);
  const int eq_take = max(0, min(cnt_eq, k - eq_before));
  int pos = block_exclusive_scan_i(cnt_sure + eq_take, warp_scan);
  int drop = i0 - pos;  // dropped entries before this chunk
  int eq_left = eq_take;
  const int64_t free_base = free_rows != nullptr ? free_rows[r] : 0;
  for (int i = i0; i < i1; ++i) {
    bool keep = i >= recent_start;
    if (!keep && heavy > 0) {
      const uint32_t key = keys[i];
      keep = key > thresh;
      if (key == thresh && eq_left > 0) {
        keep = true;
        --eq_left;
      }
    }
    if (keep) {
      if (keep_out != nullptr) keep_out[(int64_t)r * budget + pos] = i;
      cum[pos] = vals[i];
      srow[pos] = slots[i];
      ++pos;
    } else {
      if (free_stack != nullptr) free_stack[free_base + drop] = slots[i];
      ++drop;
    }
  }
  __syncthreads();
  // zero the tails (snapkv.py:1797-1803 zeroes the slot tail; scores keep `budget` entries)
  for (int i = budget + tid; i < kv_len; i += kH2oThreads) {
    srow[i] = 0;
    cum[i] = 0.f;
  }
}

int launch_h2o_normalize_accumulate(float* score, int64_t score_stride, int batch, int width,
                                    float scale, float* cumulative, int64_t cum_stride,
                                    const int32_t* kv_lens, cudaStream_t stream) {
  if (batch <= 0 || width <= 0) return 0;
  const size_t smem = (size_t)width * sizeof(float);
  if (smem <= 96 * 1024) {
    auto kern = h2o_normalize_accumulate_kernel<true>;
    if (smem > 48 * 1024) {
      cudaError_t e =
          cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return cuda_error("h2o_normalize smem attr", e);
    }
    kern<<<batch, kH2oThreads, smem, stream>>>(score, score_stride, width, scale, cumulative,
                                               cum_stride, kv_lens);
  } else {
    h2o_normalize_accumulate_kernel<false><<<batch, kH2oThreads, 0, stream>>>(
        score, score_stride, width, scale, cumulative, cum_stride, kv_lens);
  }
  return check_launch("h2o_normalize_accumulate");
}

int launch_h2o_select_compact(float* cumulative, const int64_t* cum_rows, int32_t* slot_tables,
                              const int64_t* slot_rows, int32_t* free_stack,
                              const int64_t* free_rows, int32_t* keep_out, int num_rows, int kv_len,
                              int budget, int recent, cudaStream_t stream) {
  if (num_rows <= 0) return 0;
  const size_t smem = (size_t)kv_len * 12;
  auto kern = h2o_select_compact_kernel;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_error("h2o_select_compact smem attr", e);
  }
  kern<<<num_rows, kH2oThreads, smem, stream>>>(cumulative, cum_rows, slot_tables, slot_rows,
                                               free_stack, free_rows, keep_out, kv_len, budget,
                                               recent);
  return check_launch("h2o_select_compact");
}

}  // namespace svb200
