// Experiment: what read bandwidth can B200 deliver for this path's access patterns?
// (a) sequential 16-byte streaming read, (b) random REC-byte records through an index table
// (REC = 2048: one Quest bounds record at Hkv 4; REC = 16384: one 16-token K or V page at Hkv 4).
// Each warp reads whole 512-byte lines; every thread keeps UNROLL independent 16-byte loads in flight.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/gather_ceiling scripts/exp_gather_ceiling.cu
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include <random>

template <int UNROLL>
__global__ void gather_read(const uint4* __restrict__ base, const int* __restrict__ idx, long n_rec,
                            int rec_vec /* uint4 per record */, unsigned* sink) {
  const long warps = (long)gridDim.x * (blockDim.x >> 5);
  const long w = (long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  const int lines = rec_vec / 32;  // 512-byte lines per record
  const long total = n_rec * lines;
  unsigned acc = 0;
  for (long l0 = w * UNROLL; l0 < total; l0 += warps * UNROLL) {
    uint4 v[UNROLL];
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long l = l0 + u;
      if (l < total) {
        const long rec = l / lines;
        const int li = (int)(l - rec * lines);
        const long phys = idx ? (long)idx[rec] : rec;
        asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v[u].x), "=r"(v[u].y), "=r"(v[u].z), "=r"(v[u].w)
                     : "l"(base + phys * rec_vec + li * 32 + lane));
      } else {
        v[u] = make_uint4(0, 0, 0, 0);
      }
    }
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) acc ^= v[u].x ^ v[u].y ^ v[u].z ^ v[u].w;
  }
  if (acc == 0x12345678u) *sink = acc;
}

int main() {
  const size_t pool_bytes = 4ull << 30;     // 4 GiB pool >> 126 MB L2
  const size_t read_bytes = 512ull << 20;   // bytes one launch reads (cfg3 score launch: 538 MB)
  uint4* pool;
  unsigned* sink;
  cudaMalloc(&pool, pool_bytes);
  cudaMalloc(&sink, 4);
  cudaMemset(pool, 1, pool_bytes);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0);
  cudaEventCreate(&e1);
  std::mt19937_64 rng(7);
  const int recs[] = {0, 512, 1024, 2048, 4096, 16384};
  for (int rec : recs) {
    const int rec_bytes = rec ? rec : 2048;
    const long n_pool = pool_bytes / rec_bytes;
    const long n_rec = read_bytes / rec_bytes;
    int* d_idx = nullptr;
    const int reps = 6;
    if (rec) {
      std::vector<int> perm(n_pool);
      for (long i = 0; i < n_pool; ++i) perm[i] = (int)i;
      std::shuffle(perm.begin(), perm.end(), rng);
      cudaMalloc(&d_idx, n_pool * sizeof(int));
      cudaMemcpy(d_idx, perm.data(), n_pool * sizeof(int), cudaMemcpyHostToDevice);
    }
    for (int ctas_per_sm : {2, 4, 8}) {
      float best = 1e9f;
      for (int r = 0; r < reps; ++r) {
        // each repetition reads a different window of the pool (rec == 0: a different 512 MB slab)
        const long off = (long)(r % (int)(n_pool / n_rec)) * n_rec;
        cudaEventRecord(e0);
        if (rec)
          gather_read<8><<<148 * ctas_per_sm, 256>>>(pool, d_idx + off, n_rec, rec_bytes / 16, sink);
        else
          gather_read<8><<<148 * ctas_per_sm, 256>>>(pool + off * (rec_bytes / 16), nullptr, n_rec, rec_bytes / 16, sink);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (r > 0) best = std::min(best, ms);
      }
      printf("record %6d B %s  ctas/sm %d  %.1f us  %.0f GB/s\n", rec_bytes, rec ? "random    " : "sequential", ctas_per_sm,
             best * 1e3, read_bytes / (best * 1e-3) / 1e9);
    }
    if (d_idx) cudaFree(d_idx);
  }
  cudaError_t e = cudaDeviceSynchronize();
  printf("status: %s\n", cudaGetErrorString(e));
  return e != cudaSuccess;
}
