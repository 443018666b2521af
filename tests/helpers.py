"""Test helpers shared by the GPU parity tests."""
from __future__ import annotations

import numpy as np
import torch

PAGE = 16


from oracle.cases import chunk_maps  # noqa: E402,F401


def packed_bounds(page_major: torch.Tensor, page_rec: torch.Tensor, num_chunks: int, device) -> torch.Tensor:
    """Page-major bounds [P, 2, Hkv, 128] -> this provider's chunk layout on ``device``."""
    from sparse_vllm_b200 import ops
    hkv = int(page_major.shape[2])
    out = ops.alloc_bounds(num_chunks, hkv, device)
    return ops.pack_bounds(page_major.to(device), page_rec.to(device), out)
