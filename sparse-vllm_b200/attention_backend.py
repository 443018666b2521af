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


def get_decode_workspace(context, batch_size: int, num_heads: int, num_kv_heads: int,
                         device: torch.device) -> torch.Tensor:
    """Cached scratch of the decode launch (role of layers/attention.py:15-51 ``get_decode_workspace``).

    The reference caches fp32 ``mid_o [B, Hq, ceil(len/256), D]`` + ``mid_o_logexpsum``; the stream-K kernel
    needs arrival tickets plus two partial records per SM instead, sized by svb200_workspace_bytes.  The
    buffer is allocated (zeroed) once per shape, outside graph capture, and stays valid for replays."""
    key = (int(batch_size), int(num_heads), int(num_kv_heads), str(device))
    ws = context.decode_workspace
    if ws is None or context.decode_workspace_key != key:
        ws = ops.attn_workspace(batch_size, num_heads, num_kv_heads, device)
        context.decode_workspace = ws
        context.decode_workspace_key = key
    return ws


class B200AttentionBackend:
    def __init__(self, provider: B200SparseDecodeProvider | None = None) -> None:
        self.provider = provider

    def _provider(self, q, num_heads, num_kv_heads) -> B200SparseDecodeProvider:
        if self.provider is None:
            spec = SparseDecodeAttentionSpec(int(num_heads), int(num_kv_heads), int(q.shape[-1]), q.dtype)
            self.provider = prepare_sparse_decode_attention_op(spec, device_index=q.device.index).provider
        return self.provider

    def run_decode(self, q: torch.Tensor, view: DecodeComputeView, *, mid_o: torch.Tensor | None = None,
                   mid_o_logexpsum: torch.Tensor | None = None, max_len_in_batch: int, block_seq: int = 256,
                   num_heads: int, num_kv_heads: int, gqa_block_n: int = 16, gqa_num_warps: int = 2,
                   workspace: torch.Tensor | None = None, out: torch.Tensor | None = None,
                   tp_comm=None, tp_gathered: torch.Tensor | None = None) -> torch.Tensor:
        # mid_o / mid_o_logexpsum / block_seq / gqa_* are the reference's split-K workspace and Triton launch
        # knobs; they are accepted for signature compatibility and unused (stream-K has its own scratch).
        # tp_comm / tp_gathered (extension): the launch's epilogue also gathers the head shards of all TP ranks
        # into ``tp_gathered [B, world*Hq, D]`` (parallel.PeerComm), in place of a collective after the layer.
        del mid_o, mid_o_logexpsum, block_seq, gqa_block_n, gqa_num_warps
        payload = view.payload
        if not isinstance(payload, ExplicitKVPayload):
            raise TypeError(f"B200 decode requires ExplicitKVPayload, got {type(payload).__name__}.")
        if payload.backend != "dense":
            raise RuntimeError(f"B200 decode serves payload.backend='dense' only, got {payload.backend!r}.")
        meta = view.meta
        if int(num_heads) != int(q.shape[1]) or int(num_kv_heads) != int(payload.k_cache.shape[1]):
            raise ValueError("num_heads / num_kv_heads do not match q / k_cache.")
        if workspace is None:
            workspace = get_decode_workspace(get_context(), q.shape[0], num_heads, num_kv_heads, q.device)
        o = torch.empty_like(q) if out is None else out
        return self._provider(q, num_heads, num_kv_heads).decode(
            q, payload.k_cache, payload.v_cache, meta.active_slots, meta.req_indices, meta.context_lens, out=o,
            workspace=workspace, max_len=int(max_len_in_batch), attn_score=meta.attn_score,
            comm=tp_comm, gathered=tp_gathered)


def decode_attention_layer(cache_manager, backend: B200AttentionBackend, layer_idx: int, q: torch.Tensor,
                           selection: SparseSelection, *, num_heads: int, num_kv_heads: int,
                           out: torch.Tensor | None = None, tp_comm=None,
                           tp_gathered: torch.Tensor | None = None) -> torch.Tensor:
    """Decode branch of the generic ``Attention.forward`` (layers/attention.py:162-250)."""
    context = get_context()
    one_launch = getattr(cache_manager, "decode_layer_one_launch", None)
    if one_launch is not None:
        o = one_launch(layer_idx, q, selection, num_heads=num_heads, num_kv_heads=num_kv_heads, out=out,
                       workspace=get_decode_workspace(context, q.shape[0], num_heads, num_kv_heads, q.device),
                       tp_comm=tp_comm, tp_gathered=tp_gathered)
        if o is not None:
            cache_manager.on_layer_attention_end(layer_idx)
            return o
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
    workspace = get_decode_workspace(context, q.shape[0], num_heads, num_kv_heads, q.device)
    o = backend.run_decode(q, view, max_len_in_batch=max_len_in_batch, num_heads=num_heads,
                           num_kv_heads=num_kv_heads, workspace=workspace, out=out, tp_comm=tp_comm,
                           tp_gathered=tp_gathered)
    cache_manager.on_layer_attention_end(layer_idx)
    return o
