"""Tensor-parallel plumbing of the decode path: KV-head sharding, peer-mapped memory and the one exchange step.

Reference: each TP rank owns ``Hkv / tp`` KV heads and their query heads (engine/cache_manager/base.py:253-259,
distributed/sharding.py:25-39: ``Hkv % tp == 0`` is required here, KV replication for tp > Hkv is out of scope),
its own KV cache / Quest bounds, and selects pages on its LOCAL heads only (CONTEXT.md:27-29); the only
collective of the path gathers the attention output shards (or sums after a row-parallel o_proj,
layers/linear.py:581-589).  One process per GPU, ``torch.distributed`` (NCCL on the GPU box, gloo in the
CPU tests).
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass

import torch
import torch.distributed as dist

from . import _native


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


# ----------------------------------------------------------------------------------- peer-mapped memory
class _RawCudaMemory:
    """``__cuda_array_interface__`` view of a raw device allocation (uint8), zero-copy into torch."""

    def __init__(self, ptr: int, nbytes: int) -> None:
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class PeerComm:
    """One data buffer + one sync block per rank, every rank's mapped into every process (CUDA IPC over
    NVLink): the ``svb200_comm_t`` the tensor-parallel epilogues write through (include/svb200.h).

    The reference's fused all-reduce providers set up the same kind of shared workspace through FlashInfer
    (operators/all_reduce.py:139-244, H100-only); here it is ``cudaMalloc`` + ``cudaIpcGetMemHandle`` in the
    C-ABI library and an ``all_gather_object`` of the 64-byte handles.  Every rank must issue the same
    sequence of collective launches; consecutive collectives must use different regions of ``buffer``."""

    SYNC_BYTES = 256

    def __init__(self, nbytes: int, *, group=None, device: torch.device | None = None, ll_bytes: int | None = None) -> None:
        """``ll_bytes``: staging of the push epilogue's flag-in-data protocol behind the sync block (default: twice the
        data buffer, enough when that holds two gathered tensors; 0: none, the epilogue then uses fences and flags)."""
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if self.world > _native.MAX_PEERS:
            raise ValueError(f"PeerComm serves up to {_native.MAX_PEERS} ranks of one NVLink domain, got {self.world}.")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index != torch.cuda.current_device():
            raise RuntimeError("PeerComm must be created on the current CUDA device (one process per GPU).")
        self.nbytes = (int(nbytes) + 255) // 256 * 256
        lib = _native.load()
        base = ctypes.c_void_p()
        handle = ctypes.create_string_buffer(64)
        self.ll_bytes = (2 * self.nbytes if ll_bytes is None else int(ll_bytes) + 255) // 256 * 256
        _native.check(lib.svb200_symm_alloc(self.nbytes + self.SYNC_BYTES + self.ll_bytes, ctypes.byref(base), handle),
                      "svb200_symm_alloc")
        self._base = int(base.value)
        self._peer_bases: list[int] = [0] * self.world
        self._opened: list[int] = []
        handles: list[bytes] = [handle.raw]
        if self.world > 1:
            handles = [None] * self.world  # type: ignore[list-item]
            dist.all_gather_object(handles, handle.raw, group=group)
        for r in range(self.world):
            if r == self.rank:
                self._peer_bases[r] = self._base
                continue
            ptr = ctypes.c_void_p()
            _native.check(lib.svb200_ipc_open(ctypes.create_string_buffer(handles[r], 64), ctypes.byref(ptr)),
                          f"svb200_ipc_open(rank {r})")
            self._peer_bases[r] = int(ptr.value)
            self._opened.append(int(ptr.value))
        self.buffer = torch.as_tensor(_RawCudaMemory(self._base, self.nbytes), device=self.device)
        self._struct = _native.Comm()
        self._struct.world, self._struct.rank = self.world, self.rank
        for r in range(self.world):
            self._struct.peer_buf[r] = self._peer_bases[r]
            self._struct.peer_sync[r] = self._peer_bases[r] + self.nbytes
            self._struct.peer_ll[r] = self._peer_bases[r] + self.nbytes + self.SYNC_BYTES if self.ll_bytes else None
        self._struct.ll_bytes = self.ll_bytes
        if self.world > 1:
            dist.barrier(group=group)  # every rank has mapped every buffer before anybody launches

    def struct_ptr(self):
        return ctypes.byref(self._struct)

    def view(self, offset_bytes: int, shape, dtype: torch.dtype) -> torch.Tensor:
        """A tensor inside the local data buffer (the same offset names the same region on every rank)."""
        n = 1
        for d in shape:
            n *= int(d)
        nb = n * torch.empty((), dtype=dtype).element_size()
        if offset_bytes % 16 or offset_bytes < 0 or offset_bytes + nb > self.nbytes:
            raise ValueError(f"view [{offset_bytes}, {offset_bytes + nb}) is not a 16-byte aligned range of the {self.nbytes}-byte buffer.")
        return self.buffer[offset_bytes: offset_bytes + nb].view(dtype).view(*shape)

    def offset_of(self, t: torch.Tensor) -> int:
        off = t.data_ptr() - self._base
        if off < 0 or off + t.numel() * t.element_size() > self.nbytes:
            raise ValueError("tensor does not live inside this PeerComm's local buffer.")
        return off

    def close(self) -> None:
        """Unmap the peers and free the local allocation (after a barrier: nobody may still be writing)."""
        if self._base == 0:
            return
        lib = _native.load()
        torch.cuda.synchronize()
        if self.world > 1 and dist.is_initialized():
            dist.barrier(group=self.group)
        for ptr in self._opened:
            lib.svb200_ipc_close(ptr)
        self._opened = []
        self.buffer = None
        lib.svb200_symm_free(self._base)
        self._base = 0
