// HBM layout of the Quest page bounds ("metadata") of this provider.
//
// The reference keeps metadata_cache[2(max,min)][L][P_total][Hkv][D] indexed by PHYSICAL page
// (engine/cache_manager/quest.py:96-105) and gathers the pages of a row through the page table
// before it scores them (quest.py:1870-1881).  Here the bounds are laid out for the scoring scan
// instead: a *bounds chunk* holds the records of 16 consecutive LOGICAL pages of one request row,
// so a 16-page scoring unit is ONE contiguous read (one bulk async copy per ring stage, no
// page-table indirection, no per-page copies) whatever the physical page assignment is.
//
//   bounds_layer[num_chunks][hs][16 pages][2 (max, min)][hh][128] bf16,   hh = Hkv / hs
//
// hs ("head split") = number of ring stages a unit is streamed as: the smallest divisor of Hkv that
// leaves at most 4 KV heads (32 KB) per stage.  Inside every 256-byte (page, plane, head) row the
// sixteen 16-byte segments are XOR-swizzled by (page & 7): ldmatrix reads 8 pages x 16 bytes at a
// time, and with rows a multiple of 128 bytes apart an unswizzled stage would put all 8 on the same
// banks.  A record's position is rec = chunk * 16 + page_in_chunk; the cache manager assigns a chunk
// to (row, logical page / 16) when the row's page is allocated and keeps two maps on the device:
//   unit_table[row][logical page / 16] -> chunk        (read by the scoring scan)
//   page_rec[physical page]            -> rec or -1    (read by the kernels that WRITE bounds)
#pragma once

#include <stdint.h>

#include "common.cuh"

namespace svb200 {

// ring stages per 16-page unit: smallest divisor of hkv leaving <= 4 heads per stage
__host__ __device__ inline int bounds_head_split(int hkv) {
  for (int d = 1; d < hkv; ++d)
    if (hkv % d == 0 && hkv / d <= 4) return d;
  return hkv <= 4 ? 1 : hkv;
}
__host__ __device__ inline int64_t bounds_chunk_bytes(int hkv) { return (int64_t)kPage * hkv * 2 * kRowBytes; }
__host__ __device__ inline int64_t bounds_part_bytes(int hkv, int hs) { return bounds_chunk_bytes(hkv) / hs; }

// Element offset (bf16 units) inside one layer's bounds of 16-byte segment `seg` (0..15) of
// (record rec, plane, head).
__device__ __forceinline__ int64_t bounds_seg_offset(int rec, int plane, int head, int seg, int hkv, int hs) {
  const int chunk = rec >> 4, page = rec & 15, hh = hkv / hs;
  const int part = head / hh, hl = head - part * hh;
  return (((((int64_t)chunk * hs + part) * kPage + page) * 2 + plane) * hh + hl) * kHeadDim +
         ((seg ^ (page & 7)) << 3);
}

}  // namespace svb200
