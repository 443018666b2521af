"""The Python view of the bounds-chunk layout (ops.pack_bounds / unpack_bounds) agrees with the address
formula the kernels use (csrc/bounds_layout.cuh: bounds_seg_offset), element for element.  CPU only."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import cases

PAGE = 16


def _seg_offset(rec, plane, head, seg, hkv, hs):
    """bounds_seg_offset restated (element offset inside one layer)."""
    chunk, page, hh = rec >> 4, rec & 15, hkv // hs
    part, hl = head // hh, head % hh
    return (((((chunk * hs + part) * PAGE + page) * 2 + plane) * hh + hl) * 128) + ((seg ^ (page & 7)) << 3)


@pytest.mark.parametrize("hkv", [1, 2, 3, 4, 5, 6, 8])
def test_pack_matches_kernel_addressing(hkv):
    from sparse_vllm_b200 import ops
    hs = ops.bounds_head_split(hkv)
    assert hkv % hs == 0 and hkv // hs <= 4 and all(hkv % d or hkv // d > 4 for d in range(1, hs))
    rows, max_pages, num_pages = 3, 40, 130
    pt = torch.full((rows, max_pages), -1, dtype=torch.int32)
    perm = torch.randperm(num_pages, generator=torch.Generator().manual_seed(hkv)).to(torch.int32)
    pt[0, :40], pt[1, :17], pt[2, :33] = perm[:40], perm[40:57], perm[57:90]
    ut, rec, num_chunks = cases.chunk_maps(pt, num_pages, seed=hkv)
    # unique value per element so that a misplaced element is detected
    page_major = torch.arange(num_pages * 2 * hkv * 128, dtype=torch.int32).view(num_pages, 2, hkv, 128)
    bounds = torch.full(ops.bounds_shape(num_chunks, hkv), -1, dtype=torch.int32)
    ops.pack_bounds(page_major, rec, bounds)
    flat = bounds.reshape(-1).numpy()
    gen = np.random.Generator(np.random.PCG64(1))
    assigned = np.nonzero(rec.numpy() >= 0)[0]
    assert len(assigned) == 90
    for _ in range(400):
        page = int(gen.choice(assigned))
        plane, head, seg, e = int(gen.integers(2)), int(gen.integers(hkv)), int(gen.integers(16)), int(gen.integers(8))
        off = _seg_offset(int(rec[page]), plane, head, seg, hkv, hs) + e
        assert flat[off] == int(page_major[page, plane, head, seg * 8 + e])
    # unassigned chunks / pages untouched; unpack inverts pack
    back = ops.unpack_bounds(bounds, rec)
    assert torch.equal(back[assigned], page_major[assigned])
    rest = np.setdiff1d(np.arange(num_pages), assigned)
    assert not back[rest].any()
    assert int((bounds >= 0).sum()) == 90 * 2 * hkv * 128
    # every (row, unit) has its own chunk and records of one row are contiguous per 16 pages
    used = ut[ut >= 0]
    assert len(set(used.tolist())) == len(used) == 3 + 2 + 3
