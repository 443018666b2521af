"""Boundary types and the cache-manager base of the B200 provider.

Dataclasses mirror engine/cache_manager/base.py:71-211 of the reference (same names and fields for
the explicit-KV path); ``CacheManager`` keeps the hooks the generic ``Attention`` layer and the
engine call (base.py:225-1819): ``create``, ``get_layer_kv_cache``, ``save_rope_kv_if_needed``,
``on_kv_stored``, ``build_decode_view``, ``build_decode_compute_view``, ``on_forward_end``,
``on_layer_attention_end``, ``get_decode_block_seq``, ``decode_cuda_graph_context_capacity``.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Any

import torch

from .. import ops
from ..method_registry import require_provider_method


@dataclass
class LayerBatchStates:
    slot_mapping: torch.Tensor | None = None
    context_lens: torch.Tensor | None = None
    max_context_len: int | None = None
    req_indices: torch.Tensor | None = None


@dataclass
class SparseSelection:
    """Logical token selection produced by the sparse controller for one layer."""
    kind: str
    req_indices: torch.Tensor
    context_lens: torch.Tensor
    max_context_len: int | None = None
    attn_score: torch.Tensor | None = None
    active_slots: torch.Tensor | None = None


@dataclass(frozen=True)
class AttentionViewMeta:
    active_slots: torch.Tensor
    req_indices: torch.Tensor
    context_lens: torch.Tensor
    max_context_len: int | None = None
    attn_score: torch.Tensor | None = None
    temp_slots: torch.Tensor | None = None


@dataclass(frozen=True)
class ExplicitKVPayload:
    k_cache: torch.Tensor
    v_cache: torch.Tensor
    backend: str = "dense"
    metadata: dict[str, Any] | None = None


@dataclass(frozen=True)
class DecodeComputeView:
    meta: AttentionViewMeta
    payload: ExplicitKVPayload


class CacheManager:
    """Common state: paged/slotted KV ``[2, L, slots, Hkv, D]`` bf16 resident in HBM."""

    sparse_method = ""

    # Tests of the host-side bookkeeping set this to build a manager on CPU tensors with the launches
    # replaced by recorders; every op still refuses non-CUDA tensors (ops.require_cuda).
    _allow_host_tensors = False

    def __init__(self, *, num_layers: int, num_kv_heads: int, head_dim: int, num_slots: int,
                 device: torch.device | str, dtype: torch.dtype = torch.bfloat16) -> None:
        if head_dim != ops.HEAD_DIM or dtype != torch.bfloat16:
            raise ValueError(f"the B200 provider serves bf16 KV with head_dim={ops.HEAD_DIM} only.")
        self.device = torch.device(device)
        if self.device.type != "cuda" and not self._allow_host_tensors:
            raise RuntimeError("the B200 provider has no CPU path: cache managers live on a CUDA device.")
        self.num_layers = int(num_layers)
        self.num_kv_heads = int(num_kv_heads)
        self.head_dim = int(head_dim)
        self.num_slots = int(num_slots)
        self.kv_cache = torch.zeros((2, self.num_layers, self.num_slots, self.num_kv_heads, self.head_dim),
                                    dtype=dtype, device=self.device)
        self.layer_batch_state = LayerBatchStates()
        self.kv_layer_map = None  # optional callable: logical layer -> physical KV layer

    def kv_layer_index(self, layer_idx: int) -> int:
        """Logical layer -> index into the KV storage (runtime_layout.kv_layer_index in the reference)."""
        idx = int(layer_idx) if self.kv_layer_map is None else int(self.kv_layer_map(int(layer_idx)))
        if not 0 <= idx < self.num_layers:
            raise IndexError(f"layer {layer_idx} maps to KV layer {idx}, outside [0, {self.num_layers}).")
        return idx

    # ---- factory (base.py:329-369) ---------------------------------------------------------
    @staticmethod
    def create(config, **kwargs) -> "CacheManager":
        method = require_provider_method(getattr(config, "vllm_sparse_method", ""))
        if method == "quest":
            from .quest import QuestCacheManager
            return QuestCacheManager(config, **kwargs)
        from .h2o import H2OCacheManager
        return H2OCacheManager(config, **kwargs)

    # ---- hooks ----------------------------------------------------------------------------
    def get_layer_kv_cache(self, layer_idx: int) -> tuple[torch.Tensor, torch.Tensor]:
        kv_idx = self.kv_layer_index(layer_idx)
        return self.kv_cache[0, kv_idx], self.kv_cache[1, kv_idx]

    def save_rope_kv_if_needed(self, layer_idx: int, k: torch.Tensor, v: torch.Tensor,
                               slot_mapping: torch.Tensor | None = None) -> None:
        """Persist the current tokens' K/V rows (base.py:674-694 -> store_kvcache) then ``on_kv_stored``."""
        if slot_mapping is None:
            slot_mapping = self.layer_batch_state.slot_mapping
        k_cache, v_cache = self.get_layer_kv_cache(layer_idx)
        ops.store_kv_minmax(k, v, k_cache, v_cache, slot_mapping, self._fused_store_bounds(layer_idx),
                            getattr(self, "page_rec", None))
        self.on_kv_stored(layer_idx, k, slot_mapping)

    def _fused_store_bounds(self, layer_idx: int) -> torch.Tensor | None:
        return None

    def on_kv_stored(self, layer_idx: int, k: torch.Tensor, slot_mapping: torch.Tensor) -> None:
        return None

    def on_forward_end(self, seqs, is_prefill: bool) -> None:
        return None

    def on_layer_attention_end(self, layer_idx: int) -> None:
        return None

    def get_decode_block_seq(self, layer_idx: int, default: int) -> int:
        return int(default)

    def decode_cuda_graph_context_capacity(self, max_context_len: int) -> int:
        return int(max_context_len)

    def _default_active_slots_for_selection(self, layer_idx: int, selection: SparseSelection) -> torch.Tensor:
        raise NotImplementedError

    def build_decode_view(self, layer_idx, q, active_slots, req_indices, context_lens, *, num_heads, num_kv_heads):
        return active_slots, req_indices, context_lens

    def build_decode_compute_view(self, layer_idx: int, q: torch.Tensor, selection: SparseSelection, *,
                                  num_heads: int, num_kv_heads: int) -> DecodeComputeView:
        """base.py:1162-1206."""
        active_slots = selection.active_slots
        if active_slots is None:
            active_slots = self._default_active_slots_for_selection(layer_idx, selection)
        active_slots, req_indices, context_lens = self.build_decode_view(
            layer_idx, q, active_slots, selection.req_indices, selection.context_lens,
            num_heads=num_heads, num_kv_heads=num_kv_heads)
        k_cache, v_cache = self.get_layer_kv_cache(layer_idx)
        return DecodeComputeView(
            meta=AttentionViewMeta(active_slots=active_slots, req_indices=req_indices, context_lens=context_lens,
                                   attn_score=selection.attn_score, max_context_len=selection.max_context_len),
            payload=ExplicitKVPayload(k_cache=k_cache, v_cache=v_cache))
