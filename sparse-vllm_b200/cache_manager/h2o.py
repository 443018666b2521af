"""H2O cache manager of the B200 provider: scored decode + burst eviction, state kept on the GPU.

Reference: ``H2OCacheManager`` (engine/cache_manager/h2o.py) on top of the per-layer slot tables of
``SnapKVCacheManager`` (engine/cache_manager/snapkv.py), and the H2O parts of the sparse controller
(engine/sparse_controller.py:427-497 score buffer, 545-594 reset, 761-767 layer-end normalise).

What lives where
* ``slot_tables [L, rows, W] int32``: every layer keeps its own logical -> physical token slots (layers
  evict different tokens; snapkv.py per-layer ``buffer_req_to_token_slots``).  KV bytes never move.
* ``free_stack [L, num_slots] int32`` + one host count: rows evict on the same length trigger, so every
  layer pops / pushes the same NUMBER of slots and the count is host-known (no device sync).
* ``step_scores [L, B, W] fp32``: the 2D head-max score buffer the attention kernel max-updates
  (sparse_controller.py:427-497), refilled with -1e20 every step (545-594).
* ``cumulative [L * rows, W] fp32``: persistent cumulative scores; the reference rebuilds L*B row views
  of a clone every step (h2o.py:958-1038), here one fused kernel normalises the step scores and
  accumulates in place.
"""
from __future__ import annotations

import math

import numpy as np
import torch

from .. import ops
from ..operators.decode_attention import (B200SparseDecodeProvider, SparseDecodeAttentionSpec,
                                          prepare_sparse_decode_attention_op)
from .base import CacheManager, SparseSelection

SCORE_FILL = -1e20  # sparse_controller.py:427-497


class H2OCacheManager(CacheManager):
    sparse_method = "h2o"

    def __init__(self, config, *, num_layers: int, num_kv_heads: int, num_slots: int, max_rows: int,
                 device, head_dim: int = ops.HEAD_DIM, provider: B200SparseDecodeProvider | None = None) -> None:
        super().__init__(num_layers=num_layers, num_kv_heads=num_kv_heads, head_dim=head_dim,
                         num_slots=num_slots, device=device)
        self.config = config
        self.budget = int(config.h2o_decode_budget)
        self.interval = int(config.h2o_decode_eviction_interval)
        self.recent_ratio = float(config.h2o_recent_ratio)
        if self.budget <= 0 or self.interval <= 0:
            raise ValueError("h2o_decode_budget and h2o_eviction_interval must be positive.")
        self.max_rows = int(max_rows)
        # a row never holds more than budget + interval tokens between two evictions
        self.width = self.budget + self.interval
        L, R, W = self.num_layers, self.max_rows, self.width
        self.slot_tables = torch.zeros((L, R, W), dtype=torch.int32, device=self.device)
        self.cumulative = torch.zeros((L * R, W), dtype=torch.float32, device=self.device)
        self.step_scores = torch.full((L, R, W), SCORE_FILL, dtype=torch.float32, device=self.device)
        self.free_stack = torch.arange(self.num_slots, dtype=torch.int32, device=self.device).repeat(L, 1).contiguous()
        self.num_free = self.num_slots
        self.row_seq_lens = np.zeros((R,), dtype=np.int32)
        self.seq_id_to_row: dict[int, int] = {}
        self._free_rows = list(range(R - 1, -1, -1))
        # [L, rows] slots of the current decode token; a static buffer so that a captured layer loop replays
        self._layer_slot_mapping = torch.zeros((L, R), dtype=torch.int32, device=self.device)
        self._batch = 0
        self._kv_lens_dev = torch.zeros((R,), dtype=torch.int32, device=self.device)
        self._rows_dev: torch.Tensor | None = None
        self._batch_rows: list[int] = []
        self._provider = provider

    # ------------------------------------------------------------------ provider
    def _get_provider(self, num_heads: int) -> B200SparseDecodeProvider:
        if self._provider is None:
            spec = SparseDecodeAttentionSpec(int(num_heads), self.num_kv_heads, self.head_dim, torch.bfloat16,
                                             page_size=1, sparse_method="h2o")
            self._provider = prepare_sparse_decode_attention_op(spec, device_index=self.device.index).provider
        return self._provider

    # ------------------------------------------------------------------ allocation
    @property
    def num_free_slots(self) -> int:
        return self.num_free

    def _get_free_row(self, seq_id: int) -> int:
        row = self.seq_id_to_row.get(seq_id)
        if row is None:
            if not self._free_rows:
                raise RuntimeError("Out of H2O request rows.")
            row = self._free_rows.pop()
            self.seq_id_to_row[seq_id] = row
            self.row_seq_lens[row] = 0
        return row

    def _pop_slots(self, n: int) -> torch.Tensor:
        """Top ``n`` entries of every layer's free stack -> [L, n] (the count is common to all layers)."""
        if self.num_free < n:
            raise RuntimeError(f"Out of H2O KV slots: need={n}, free={self.num_free}.")
        out = self.free_stack[:, self.num_free - n: self.num_free]
        self.num_free -= n
        return out

    def admit_compacted_prompt(self, seq_id: int, k: torch.Tensor, v: torch.Tensor, scores: torch.Tensor) -> int:
        """Admit a sequence whose prompt was already compacted to <= budget tokens per layer by the prefill
        side (h2o.py:1181-1350, out of this provider's scope): k, v ``[L, T, Hkv, D]`` bf16 and the kept
        tokens' cumulative scores ``[L, T]`` fp32.  Returns the row."""
        L, T = int(k.shape[0]), int(k.shape[1])
        if L != self.num_layers or T > self.budget:
            raise ValueError(f"expected [L={self.num_layers}, T<=budget={self.budget}, Hkv, D], got {tuple(k.shape)}.")
        row = self._get_free_row(seq_id)
        slots = self._pop_slots(T)
        self.slot_tables[:, row, :T] = slots
        self.cumulative.view(L, self.max_rows, self.width)[:, row, :T] = scores
        for layer in range(L):
            k_cache, v_cache = self.get_layer_kv_cache(layer)
            ops.store_kv_minmax(k[layer], v[layer], k_cache, v_cache, slots[layer].contiguous())
        self.row_seq_lens[row] = T
        return row

    def free_seq(self, seq_id: int) -> None:
        row = self.seq_id_to_row.pop(seq_id, None)
        if row is None:
            return
        n = int(self.row_seq_lens[row])
        if n:
            self.free_stack[:, self.num_free: self.num_free + n] = self.slot_tables[:, row, :n]
            self.num_free += n
        self.row_seq_lens[row] = 0
        self._free_rows.append(row)

    # ------------------------------------------------------------------ decode step (host)
    def prepare_decode_static(self, seqs, input_ids, positions, slot_mapping, context_lens, req_indices) -> None:
        """One slot per row AND layer for the token being decoded, static metadata buffers filled in place
        (quest.py:1280-1360 pattern; H2O slots are per layer, see ``layer_slot_mapping``)."""
        bsz = len(seqs)
        rows = [self.seq_id_to_row[int(s.seq_id)] for s in seqs]
        lens = self.row_seq_lens[rows]
        if int(lens.max()) + 1 > self.width:
            raise RuntimeError("row longer than budget + interval: evict_after_decode() was not called.")
        new = self._pop_slots(bsz)  # [L, B]
        rows_dev = torch.as_tensor(rows, dtype=torch.int64, device=self.device)
        lens_dev = torch.as_tensor(lens.astype(np.int64), device=self.device)
        self.slot_tables[:, rows_dev, lens_dev] = new
        self._layer_slot_mapping[:, :bsz] = new
        self._batch = bsz
        self.row_seq_lens[rows] = lens + 1
        ctx = torch.as_tensor((lens + 1).astype(np.int32), device=self.device)
        context_lens[:bsz] = ctx
        req_indices[:bsz] = rows_dev.to(torch.int32)
        slot_mapping[:bsz] = new[0]
        self._kv_lens_dev[:bsz] = ctx
        self._rows_dev = rows_dev
        self._batch_rows = rows
        self.layer_batch_state.slot_mapping = slot_mapping[:bsz]
        self.layer_batch_state.context_lens = context_lens[:bsz]
        self.layer_batch_state.req_indices = req_indices[:bsz]
        self.layer_batch_state.max_context_len = self.width
        self.step_scores[:, :bsz].fill_(SCORE_FILL)  # sparse_controller.py:545-594

    def layer_slot_mapping(self, layer_idx: int) -> torch.Tensor:
        return self._layer_slot_mapping[self.kv_layer_index(layer_idx), : self._batch]

    def save_rope_kv_if_needed(self, layer_idx, k, v, slot_mapping=None) -> None:
        k_cache, v_cache = self.get_layer_kv_cache(layer_idx)
        ops.store_kv_minmax(k, v, k_cache, v_cache,
                            self.layer_slot_mapping(layer_idx) if slot_mapping is None else slot_mapping)

    # ------------------------------------------------------------------ per-layer hooks
    def decode_selection(self, layer_idx: int) -> SparseSelection:
        """Everything the row holds is attended; the kernel max-updates the layer's 2D score rows."""
        st = self.layer_batch_state
        bsz = int(st.context_lens.numel())
        kv_idx = self.kv_layer_index(layer_idx)
        return SparseSelection(kind="h2o", req_indices=st.req_indices, context_lens=st.context_lens,
                               max_context_len=self.width, attn_score=self.step_scores[kv_idx, :bsz],
                               active_slots=self.slot_tables[kv_idx])

    def _default_active_slots_for_selection(self, layer_idx, selection):
        return self.slot_tables[self.kv_layer_index(layer_idx)]

    def on_layer_attention_end(self, layer_idx: int) -> None:
        """softmax(score * D^-0.5) in place (sparse_controller.py:761-767) fused with the cumulative update
        (h2o.py:958-1038): cum[:kv_len-1] += step, cum[kv_len-1] = step."""
        kv_idx = self.kv_layer_index(layer_idx)
        bsz = len(self._batch_rows)
        rows = self._batch_rows
        cum = self.cumulative.view(self.num_layers, self.max_rows, self.width)[kv_idx]
        if rows != list(range(rows[0], rows[0] + bsz)):
            raise NotImplementedError("decode batches must use consecutive rows (the fused update indexes them by stride).")
        ops.h2o_normalize_accumulate(self.step_scores[kv_idx, :bsz], 1.0 / math.sqrt(self.head_dim),
                                     cumulative=cum[rows[0]: rows[0] + bsz], kv_lens=self._kv_lens_dev[:bsz],
                                     width=self.width)

    # ------------------------------------------------------------------ eviction
    def recent_count(self, kv_len: int) -> int:
        return min(max(1, int(self.budget * self.recent_ratio)), self.budget, kv_len)  # h2o.py:498-499

    def evict_after_decode(self, keep_out: torch.Tensor | None = None) -> bool:
        """Burst eviction (h2o.py:1498-1625): when the batch rows reached budget + interval tokens (or
        the free stack is empty and they exceed the budget) every (layer, row) keeps the heavy hitters
        plus the recent suffix; slot rows and cumulative scores are compacted in place, dropped slots go
        back to the layer's free stack.  Returns True if an eviction ran."""
        rows = self._batch_rows
        if not rows:
            return False
        lens = self.row_seq_lens[rows]
        kv_len = int(lens[0])
        if not (lens == kv_len).all():
            raise NotImplementedError("rows of one decode batch evict on a common length.")
        if not (kv_len >= self.budget + self.interval or (self.num_free < len(rows) and kv_len >= self.budget + 1)):
            return False
        L, R, W = self.num_layers, self.max_rows, self.width
        drop = kv_len - self.budget
        lr = (torch.arange(L, device=self.device, dtype=torch.int64)[:, None] * R + self._rows_dev[None, :]).reshape(-1)
        offs = lr * W
        within = torch.arange(len(rows), device=self.device, dtype=torch.int64) * drop
        free_rows = (torch.arange(L, device=self.device, dtype=torch.int64)[:, None] * self.num_slots
                     + self.num_free + within[None, :]).reshape(-1)
        ops.h2o_select_compact(self.cumulative, offs, self.slot_tables, offs, kv_len=kv_len, budget=self.budget,
                               recent=self.recent_count(kv_len), free_stack=self.free_stack, free_rows=free_rows,
                               keep_out=keep_out)
        self.num_free += len(rows) * drop
        self.row_seq_lens[rows] = self.budget
        return True
