"""Decode attention backend and the generic decode layer of the B200 provider.

``B200AttentionBackend.run_decode`` has the signature of ``TritonAttentionBackend.run_decode``
(layers/attention_backend.py:217-230) and replaces its stage-1(+score) / stage-2 pair with one
sm_100a launch; ``decode_attention_layer`` restates the decode branch of ``Attention.forward``
(layers/attention.py:162-250), which stays generic: it only talks to cache-manager hooks and the backend.
"""
from __future__ import annotations

import torch

from . import ops
from .cache_manager.base import DecodeComputeView, ExplicitKVPayload, SparseSelection
from .operators.decode_attention import (B200SparseDecodeProvider, SparseDecodeAttentionSpec,
                                         prepare_sparse_decode_attention_op)
from .utils_context import get_context


def get_decode_workspace(context, batch_size: int, num_heads: int, num_blocks: int, head_dim: int,
                         device: torch.device, num_kv_heads: int | None = None):
    """Cached fp32 split-K workspace, handed out as sliced views (layers/attention.py:15-51), plus the
    zero-initialised arrival tickets of the fused combine."""
    mid_o = context.decode_mid_o
    if (mid_o is None or mid_o.device != device or mid_o.shape[0] < batch_size or mid_o.shape[1] < num_heads
            or mid_o.shape[2] < num_blocks or mid_o.shape[3] < head_dim):
        mid_o = torch.empty((batch_size, num_heads, num_blocks, head_dim), dtype=torch.float32, device=device)
        context.decode_mid_o = mid_o
    mid_lse = context.decode_mid_o_logexpsum
    if (mid_lse is None or mid_lse.device != device or mid_lse.shape[0] < batch_size
            or mid_lse.shape[1] < num_heads or mid_lse.shape[2] < num_blocks):
        mid_lse = torch.empty((batch_size, num_heads, num_blocks), dtype=torch.float32, device=device)
        context.decode_mid_o_logexpsum = mid_lse
    need = batch_size * (num_kv_heads if num_kv_heads is not None else num_heads)
    tickets = context.decode_tickets
    if tickets is None or tickets.device != device or tickets.numel() < need:
        tickets = torch.zeros((need,), dtype=torch.int32, device=device)
        context.decode_tickets = tickets
    return (mid_o[:batch_size, :num_heads, :num_blocks, :head_dim], mid_lse[:batch_size, :num_heads, :num_blocks],
            tickets)


class B200AttentionBackend:
    def __init__(self, provider: B200SparseDecodeProvider | None = None) -> None:
        self.provider = provider

    def _provider(self, q, num_heads, num_kv_heads) -> B200SparseDecodeProvider:
        if self.provider is None:
            spec = SparseDecodeAttentionSpec(int(num_heads), int(num_kv_heads), int(q.shape[-1]), q.dtype)
            self.provider = prepare_sparse_decode_attention_op(spec, device_index=q.device.index).provider
        return self.provider

    def run_decode(self, q: torch.Tensor, view: DecodeComputeView, *, mid_o: torch.Tensor,
                   mid_o_logexpsum: torch.Tensor, max_len_in_batch: int, block_seq: int, num_heads: int,
                   num_kv_heads: int, gqa_block_n: int = 16, gqa_num_warps: int = 2,
                   tickets: torch.Tensor | None = None, out: torch.Tensor | None = None) -> torch.Tensor:
        del gqa_block_n, gqa_num_warps  # Triton launch knobs of the reference; meaningless here
        payload = view.payload
        if not isinstance(payload, ExplicitKVPayload):
            raise TypeError(f"B200 decode requires ExplicitKVPayload, got {type(payload).__name__}.")
        if payload.backend != "dense":
            raise RuntimeError(f"B200 decode serves payload.backend='dense' only, got {payload.backend!r}.")
        meta = view.meta
        if int(num_heads) != int(q.shape[1]) or int(num_kv_heads) != int(payload.k_cache.shape[1]):
            raise ValueError("num_heads / num_kv_heads do not match q / k_cache.")
        if tickets is None:
            ctx = get_context()
            need = int(q.shape[0]) * int(num_kv_heads)
            if ctx.decode_tickets is None or ctx.decode_tickets.device != q.device or ctx.decode_tickets.numel() < need:
                ctx.decode_tickets = torch.zeros((need,), dtype=torch.int32, device=q.device)
            tickets = ctx.decode_tickets
        o = torch.empty_like(q) if out is None else out
        return self._provider(q, num_heads, num_kv_heads).decode(
            q, payload.k_cache, payload.v_cache, meta.active_slots, meta.req_indices, meta.context_lens, out=o,
            mid_o=mid_o, mid_lse=mid_o_logexpsum, tickets=tickets, max_len=int(max_len_in_batch),
            block_seq=int(block_seq), attn_score=meta.attn_score)


def decode_attention_layer(cache_manager, backend: B200AttentionBackend, layer_idx: int, q: torch.Tensor,
                           selection: SparseSelection, *, num_heads: int, num_kv_heads: int,
                           out: torch.Tensor | None = None) -> torch.Tensor:
    """Decode branch of the generic ``Attention.forward`` (layers/attention.py:162-250)."""
    context = get_context()
    view = cache_manager.build_decode_compute_view(layer_idx, q, selection, num_heads=num_heads,
                                                   num_kv_heads=num_kv_heads)
    meta = view.meta
    max_context_len = meta.max_context_len
    static_cap = getattr(cache_manager, "_decode_static_max_context_len", None)
    if static_cap is not None:
        max_context_len = max(int(max_context_len) if max_context_len is not None else 0, int(static_cap))
    if max_context_len is None:
        raise RuntimeError(f"static decode requires max_context_len, got None at layer={layer_idx}")
    max_len_in_batch = int(max_context_len)
    if meta.active_slots.dim() == 2:
        max_len_in_batch = min(max_len_in_batch, int(meta.active_slots.shape[1]))
    if max_len_in_batch <= 0:
        raise RuntimeError(f"decode requires a positive context length, got {max_len_in_batch} at layer={layer_idx}")
    block_seq = cache_manager.get_decode_block_seq(layer_idx, 256)
    num_blocks = (max_len_in_batch + block_seq - 1) // block_seq
    mid_o, mid_lse, tickets = get_decode_workspace(context, q.shape[0], num_heads, num_blocks, q.shape[-1], q.device,
                                                   num_kv_heads)
    o = backend.run_decode(q, view, mid_o=mid_o, mid_o_logexpsum=mid_lse, max_len_in_batch=max_len_in_batch,
                           block_seq=block_seq, num_heads=num_heads, num_kv_heads=num_kv_heads, tickets=tickets,
                           out=out)
    cache_manager.on_layer_attention_end(layer_idx)
    return o
