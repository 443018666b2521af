"""Tensor-parallel plumbing of the decode path: KV-head sharding and the one exchange step.

Reference: each TP rank owns ``Hkv / tp`` KV heads and their query heads (engine/cache_manager/base.py:253-259,
distributed/sharding.py:25-39: ``Hkv % tp == 0`` is required here, KV replication for tp > Hkv is out of scope),
its own KV cache / Quest bounds, and selects pages on its LOCAL heads only (CONTEXT.md:27-29); the only
collective of the path gathers the attention output shards (or sums after a row-parallel o_proj,
layers/linear.py:581-589).  One process per GPU, ``torch.distributed`` (NCCL on the GPU box, gloo in the
CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass(frozen=True)
class HeadShard:
    rank: int
    world: int
    q_heads: int    # local query heads
    kv_heads: int   # local KV heads
    q_begin: int    # first global query head of this rank
    kv_begin: int   # first global KV head of this rank

    def q_slice(self) -> slice:
        return slice(self.q_begin, self.q_begin + self.q_heads)

    def kv_slice(self) -> slice:
        return slice(self.kv_begin, self.kv_begin + self.kv_heads)


def shard_heads(num_q_heads: int, num_kv_heads: int, world: int, rank: int) -> HeadShard:
    """Contiguous head ranges per rank: KV heads [r*Hkv/tp, (r+1)*Hkv/tp) and their query groups."""
    if world <= 0 or not 0 <= rank < world:
        raise ValueError(f"rank={rank} outside world={world}.")
    if num_q_heads <= 0 or num_kv_heads <= 0 or num_q_heads % num_kv_heads:
        raise ValueError(f"num_q_heads={num_q_heads} must be a positive multiple of num_kv_heads={num_kv_heads}.")
    if num_kv_heads % world:
        raise ValueError(f"{num_kv_heads} KV heads cannot be sharded over {world} ranks "
                         "(KV-head replication for tp > Hkv is not provided).")
    kv = num_kv_heads // world
    group = num_q_heads // num_kv_heads
    return HeadShard(rank, world, kv * group, kv, rank * kv * group, rank * kv)


def gather_attention_output(o_local: torch.Tensor, *, group=None, out: torch.Tensor | None = None) -> torch.Tensor:
    """All-gather the ``[B, Hq/tp, D]`` shards.  Returns ``[tp, B, Hq/tp, D]`` (rank-major, what one
    ``all_gather_into_tensor`` writes; ``heads_first`` gives the ``[B, Hq, D]`` view order o_proj reads)."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    o_local = o_local.contiguous()
    if out is None:
        out = torch.empty((world, *o_local.shape), dtype=o_local.dtype, device=o_local.device)
    if world == 1:
        out.view(-1).copy_(o_local.view(-1))
        return out
    if dist.get_backend(group) == "gloo":
        dist.all_gather(list(out.unbind(0)), o_local, group=group)
    else:
        dist.all_gather_into_tensor(out.view(-1), o_local.view(-1), group=group)
    return out


def heads_first(gathered: torch.Tensor) -> torch.Tensor:
    """``[tp, B, Hq/tp, D]`` -> ``[B, Hq, D]`` (a strided view; rank r's heads land at r*Hq/tp ...)."""
    world, bsz, hq_local, d = gathered.shape
    return gathered.permute(1, 0, 2, 3).reshape(bsz, world * hq_local, d)
