// Tensor-parallel collectives of the decode path over NVLink peer memory (sm_100a).
//
// Reference behaviour (CURRENTF/Sparse-vLLM): RowParallelLinear.forward all-reduces [B, hidden] after a
// row-parallel o_proj (layers/linear.py:581-589 -> distributed/parallel_context.py:87-95, NCCL), captured in
// the decode graph; its fused providers (operators/all_reduce.py:115-409, FlashInfer) are gated to H100.
// Here: a one-shot all-reduce over peer-mapped buffers (every rank reads every rank's input through NVLink
// and sums in rank order, so all ranks get bit-identical results) — latency-bound messages (cfg5: 160 KB)
// need one launch and one flag round trip instead of NCCL's protocol.  The all-gather of the attention
// head shards is not a kernel of its own: decode_attn.cu pushes its output rows into the peers' buffers.
//
// Synchronisation: every rank owns a sync block int32[16] (kSync* in launch.h) inside peer-mapped memory.
// A launch takes epoch e = local epoch + 1, raises flag[rank] = e on every peer once its input is ready,
// and waits until all local flags reached e.  Flags only grow; a peer can be at most one collective ahead,
// so callers alternate two input buffers.
#include <cuda_bf16.h>

#include "common.cuh"
#include "launch.h"

namespace svb200 {

constexpr int kArThreads = 256;

__device__ __forceinline__ uint4 ld_peer_16(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}

__global__ void __launch_bounds__(kArThreads) allreduce_sum_bf16_kernel(AllReduceArgs a) {
  pdl_wait();
  int32_t* local = a.sync[a.rank];
  __shared__ int s_epoch;
  // the epoch only changes when the LAST CTA of a launch finishes: every CTA of this launch reads the same value
  if (threadIdx.x == 0) s_epoch = *reinterpret_cast<volatile int32_t*>(local + kSyncEpoch) + 1;
  __syncthreads();
  const int32_t e = s_epoch;
  const int peer = (int)threadIdx.x;
  // this rank's input was written by earlier work on the stream: announce it, then wait for everybody's
  if (blockIdx.x == 0 && peer < a.world && peer != a.rank) st_release_sys(a.sync[peer] + a.rank, e);
  if (peer < a.world && peer != a.rank) {
    SpinGuard guard;
    while (ld_acquire_sys(local + peer) - e < 0) {
      __nanosleep(64);
      guard.tick(kPeerSpinLimitNs);
    }
  }
  __syncthreads();
  pdl_launch_dependents();
  const int64_t nvec = a.n / 8;
  for (int64_t v = (int64_t)blockIdx.x * kArThreads + threadIdx.x; v < nvec; v += (int64_t)gridDim.x * kArThreads) {
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = 0.f;
    uint4 x[kMaxPeers];
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r)
      if (r < a.world) x[r] = ld_peer_16(reinterpret_cast<const uint4*>(a.in[r]) + v);
#pragma unroll
    for (int r = 0; r < kMaxPeers; ++r) {  // fixed rank order: every rank computes the same sum
      if (r < a.world) {
        const uint32_t w[4] = {x[r].x, x[r].y, x[r].z, x[r].w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          acc[2 * i] += bf16_bits_to_f32(w[i] & 0xffffu);
          acc[2 * i + 1] += bf16_bits_to_f32(w[i] >> 16);
        }
      }
    }
    uint4 o;
    o.x = pack_bf16x2(acc[0], acc[1]);
    o.y = pack_bf16x2(acc[2], acc[3]);
    o.z = pack_bf16x2(acc[4], acc[5]);
    o.w = pack_bf16x2(acc[6], acc[7]);
    reinterpret_cast<uint4*>(a.out)[v] = o;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    if (atomicAdd(local + kSyncDone, 1) == (int)gridDim.x - 1) {
      local[kSyncDone] = 0;
      local[kSyncEpoch] = e;
    }
  }
}

int launch_allreduce_sum_bf16(const AllReduceArgs& a, cudaStream_t stream) {
  if (a.n <= 0) return 0;
  const int64_t nvec = a.n / 8;
  int grid = (int)((nvec + kArThreads - 1) / kArThreads);
  if (grid > 64) grid = 64;  // a few CTAs saturate the links for latency-bound messages
  if (grid < 1) grid = 1;
  const cudaError_t e = launch_kernel(allreduce_sum_bf16_kernel, dim3(grid), dim3(kArThreads), 0, stream, a);
  if (e != cudaSuccess) return cuda_error("allreduce_sum_bf16 launch", e);
  return check_launch("allreduce_sum_bf16");
}

}  // namespace svb200
