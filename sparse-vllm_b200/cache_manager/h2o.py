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
                 device, head_dim: int = ops.HEAD_DIM, provider: B200SparseDecodeProvider | None = None,
                 max_prefill_chunk: int = 0) -> None:
        super().__init__(num_layers=num_layers, num_kv_heads=num_kv_heads, head_dim=head_dim,
                         num_slots=num_slots, device=device)
        self.config = config
        self.budget = int(config.h2o_decode_budget)
        self.interval = int(config.h2o_decode_eviction_interval)
        self.recent_ratio = float(config.h2o_recent_ratio)
        if self.budget <= 0 or self.interval <= 0:
            raise ValueError("h2o_decode_budget and h2o_eviction_interval must be positive.")
        self.max_rows = int(max_rows)
        self.prefill_budget = int(getattr(config, "h2o_prefill_budget", self.budget))
        self.score_window = int(getattr(config, "h2o_prefill_score_window", 128))
        # a row never holds more than budget + interval tokens between two decode evictions, and during chunked
        # prefill at most the prefill budget plus one chunk (h2o.py:665-748)
        self.max_prefill_chunk = int(max_prefill_chunk)
        self.decode_width = self.budget + self.interval
        self.width = max(self.decode_width,
                         self.prefill_budget + self.max_prefill_chunk if self.max_prefill_chunk > 0 else 0)
        L, R, W = self.num_layers, self.max_rows, self.width
        self.slot_tables = torch.zeros((L, R, W), dtype=torch.int32, device=self.device)
        self.cumulative = torch.zeros((L * R, W), dtype=torch.float32, device=self.device)
        self.step_scores = torch.full((L, R, self.decode_width), SCORE_FILL, dtype=torch.float32, device=self.device)
        self.free_stack = torch.arange(self.num_slots, dtype=torch.int32, device=self.device).repeat(L, 1).contiguous()
        self.num_free = self.num_slots
        self.row_seq_lens = np.zeros((R,), dtype=np.int32)
        self.seq_id_to_row: dict[int, int] = {}
        self._free_rows = list(range(R - 1, -1, -1))
        # [L, rows] slots of the current decode token; a static buffer so that a captured layer loop replays
        self._layer_slot_mapping = torch.zeros((L, R), dtype=torch.int32, device=self.device)
        self._batch = 0
        self._kv_lens_dev = torch.zeros((R,), dtype=torch.int32, device=self.device)
        self._prep_host: list[torch.Tensor] | None = None
        self._prep_dev: torch.Tensor | None = None
        self._rows_dev_i32 = torch.zeros((R,), dtype=torch.int32, device=self.device)
        self._batch_rows: list[int] = []
        # rows that decoded since they were admitted (h2o.py:1498-1516 `_h2o_active_decode_seq_ids`): under slot
        # pressure every such row that is over budget is reclaimable, scheduled in this step or not
        self._active_decode_rows: set[int] = set()
        # the layer-end kernel hands the step-score rows back pre-filled with SCORE_FILL (no refill launch per step)
        self.refill_in_kernel = True
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
        self._active_decode_rows.discard(row)
        self._free_rows.append(row)

    # ------------------------------------------------------------------ prefill side
    def allocate_prefill_chunk(self, seq_id: int, num_tokens: int) -> torch.Tensor:
        """Append a chunk of ``num_tokens`` prompt tokens to the sequence's compressed physical row in EVERY
        layer (``_prepare_prefill``, h2o.py:665-748): returns the ``[L, num_tokens]`` slots to store the chunk's
        K/V rows at (``save_rope_kv_if_needed(layer, k, v, slots[layer])``)."""
        row = self._get_free_row(seq_id)
        cur = int(self.row_seq_lens[row])
        n = int(num_tokens)
        if cur + n > self.width:
            raise RuntimeError(f"H2O prefill chunk does not fit the row: have={cur} chunk={n} width={self.width} "
                               "(construct the manager with max_prefill_chunk >= the chunk size).")
        slots = self._pop_slots(n)
        self.slot_tables[:, row, cur: cur + n] = slots
        self.row_seq_lens[row] = cur + n
        return slots

    def collect_prefill_attention_score(self, layer_idx: int, q: torch.Tensor, seq_ids: list[int],
                                        chunk_lens: list[int], b_start_loc: torch.Tensor | None = None) -> torch.Tensor:
        """Score the layer's chunk (h2o.py:778-894): the last ``h2o_prefill_score_window`` queries of every
        row's chunk attend causally to the whole physical row; the max-over-heads token probabilities are added,
        weighted by the number of scoring queries, to the row's cumulative scores (new positions start at 0).
        ``q [sum(chunk_lens), Hq, D]`` holds the chunk tokens of all rows; the chunk's K rows are already stored
        (``allocate_prefill_chunk`` + ``save_rope_kv_if_needed``).  Returns the step scores ``[B, max ctx]``."""
        kv_idx = self.kv_layer_index(layer_idx)
        rows = [self.seq_id_to_row[int(sid)] for sid in seq_ids]
        ctx = [int(self.row_seq_lens[r]) for r in rows]
        cache = [c - int(n) for c, n in zip(ctx, chunk_lens)]
        if min(cache) < 0:
            raise RuntimeError(f"H2O current chunk exceeds its physical row: context={ctx} chunk={list(chunk_lens)}.")
        starts = [max(c, e - self.score_window) for c, e in zip(cache, ctx)]  # prefill_score_ranges, h2o.py:749-776
        host = np.empty((6, len(rows)), dtype=np.int32)
        host[0], host[1], host[2], host[3], host[4] = rows, ctx, cache, starts, ctx
        host[5] = np.concatenate(([0], np.cumsum(chunk_lens)[:-1]))
        meta = torch.from_numpy(host).to(self.device)
        start_loc = meta[5] if b_start_loc is None else b_start_loc
        max_ctx = max(ctx)
        key = (len(rows), max_ctx)
        if getattr(self, "_prefill_score_buf", None) is None or self._prefill_score_key != key:
            self._prefill_score_buf = torch.empty((len(rows), (max_ctx + 63) // 64 * 64), dtype=torch.float32, device=self.device)
            self._prefill_score_key = key
        k_cache, _ = self.get_layer_kv_cache(layer_idx)
        cum = self.cumulative.view(self.num_layers, self.max_rows, self.width)[kv_idx]
        ops.h2o_prefill_score(q, k_cache, self.slot_tables[kv_idx], meta[0], start_loc, meta[1], meta[2], meta[3], meta[4],
                              self._prefill_score_buf, max_query_len=max(e - s for s, e in zip(starts, ctx)),
                              max_context_len=max_ctx, cumulative=cum, cum_rows=meta[0])
        return self._prefill_score_buf[:, :max_ctx]

    def evict_after_prefill(self, seq_ids: list[int], final: bool, keep_out: torch.Tensor | None = None) -> bool:
        """After a chunk's forward (h2o.py:1352-1496): rows longer than the prefill budget (``final``: the decode
        budget) keep their heavy hitters plus the recent suffix in every layer.  Index-only compaction of slot rows
        and cumulative scores, dropped slots back to the free stacks.  (The reference additionally MOVES the kept
        K/V rows of a final chunk into the row's lowest slots, h2o.py:1181-1350; here attention always reads
        through the slot table, so the bytes stay where they are.)"""
        budget = self.budget if final else self.prefill_budget
        rows = [self.seq_id_to_row[int(sid)] for sid in seq_ids]
        evict = [r for r in rows if int(self.row_seq_lens[r]) > budget]
        self.last_evicted_rows = evict
        if not evict:
            return False
        self._compact_rows(evict, budget, keep_out)
        return True

    def _compact_rows(self, evict: list[int], budget: int, keep_out: torch.Tensor | None) -> None:
        L, R, W = self.num_layers, self.max_rows, self.width
        lens = self.row_seq_lens[evict].astype(np.int64)
        drops = lens - budget
        if self.num_free + int(drops.sum()) > self.num_slots:
            raise RuntimeError(f"H2O eviction would overflow the free-slot stack: end={self.num_free + int(drops.sum())} "
                               f"capacity={self.num_slots}.")
        rows_t = torch.as_tensor(evict, dtype=torch.int64, device=self.device)
        layer_t = torch.arange(L, device=self.device, dtype=torch.int64)[:, None]
        offs = ((layer_t * R + rows_t[None, :]) * W).reshape(-1)
        within = torch.as_tensor(np.concatenate(([0], np.cumsum(drops)[:-1])), dtype=torch.int64, device=self.device)
        free_rows = (layer_t * self.num_slots + self.num_free + within[None, :]).reshape(-1)
        kv_lens = torch.as_tensor(lens.astype(np.int32), device=self.device).repeat(L)
        max_len = int(lens.max())
        recent = min(max(1, int(budget * self.recent_ratio)), budget)  # h2o.py:498-499
        ops.h2o_select_compact(self.cumulative, offs, self.slot_tables, offs, kv_len=max_len, budget=budget,
                               recent=recent, free_stack=self.free_stack, free_rows=free_rows,
                               keep_out=keep_out, kv_lens=kv_lens)
        self.num_free += int(drops.sum())
        self.row_seq_lens[evict] = budget

    # ------------------------------------------------------------------ decode step (host)
    def prepare_decode_static(self, seqs, input_ids, positions, slot_mapping, context_lens, req_indices) -> None:
        """One slot per row AND layer for the token being decoded, static metadata buffers filled in place
        (quest.py:1280-1360 pattern; H2O slots are per layer, see ``layer_slot_mapping``)."""
        bsz = len(seqs)
        rows = [self.seq_id_to_row[int(s.seq_id)] for s in seqs]
        lens = self.row_seq_lens[rows]
        if int(lens.max()) + 1 > self.decode_width:
            raise RuntimeError("row longer than budget + interval: evict_after_decode() was not called.")
        if self.num_free < bsz:
            raise RuntimeError(f"Out of H2O KV slots: need={bsz}, free={self.num_free}.")
        # host -> device: one pinned copy of [rows, new lens, free top], one launch that pops the slots of every
        # layer, appends them to the slot rows and fills the static step buffers (the reference issues ~10 indexing
        # ops and copies here, h2o.py:257-476)
        if self._prep_host is None:  # two pinned staging buffers: a buffer is rewritten only after its copy completed
            on_gpu = self.device.type == "cuda"
            mk = lambda: torch.empty(2 * self.max_rows + 1, dtype=torch.int32)
            self._prep_host = [mk().pin_memory() if on_gpu else mk() for _ in range(2)]
            self._prep_done = [torch.cuda.Event() if on_gpu else None for _ in range(2)]
            self._prep_dev = torch.empty(2 * self.max_rows + 1, dtype=torch.int32, device=self.device)
            self._prep_turn = 0
        turn = self._prep_turn
        self._prep_turn ^= 1
        if self._prep_done[turn] is not None:
            self._prep_done[turn].synchronize()
        host = self._prep_host[turn].numpy()
        host[:bsz] = rows
        host[bsz: 2 * bsz] = lens + 1
        host[2 * bsz] = self.num_free
        self._prep_dev[: 2 * bsz + 1].copy_(self._prep_host[turn][: 2 * bsz + 1], non_blocking=True)
        if self._prep_done[turn] is not None:
            self._prep_done[turn].record()
        ops.h2o_prepare_decode(self._prep_dev, self.free_stack, self.slot_tables, self._layer_slot_mapping,
                               context_lens, req_indices, slot_mapping, self._kv_lens_dev, self._rows_dev_i32, bsz)
        self.num_free -= bsz
        self._batch = bsz
        self.row_seq_lens[rows] = lens + 1
        self._batch_rows = rows
        self.layer_batch_state.slot_mapping = slot_mapping[:bsz]
        self.layer_batch_state.context_lens = context_lens[:bsz]
        self.layer_batch_state.req_indices = req_indices[:bsz]
        self.layer_batch_state.max_context_len = self.decode_width
        if not self.refill_in_kernel:
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
                               max_context_len=self.decode_width, attn_score=self.step_scores[kv_idx, :bsz],
                               active_slots=self.slot_tables[kv_idx])

    def _default_active_slots_for_selection(self, layer_idx, selection):
        return self.slot_tables[self.kv_layer_index(layer_idx)]

    def on_layer_attention_end(self, layer_idx: int) -> None:
        """softmax(score * D^-0.5) in place (sparse_controller.py:761-767) fused with the cumulative update
        (h2o.py:958-1038): cum[:kv_len-1] += step, cum[kv_len-1] = step.  Batch entry b updates the cumulative
        row of ITS request row (any row set, in any order)."""
        kv_idx = self.kv_layer_index(layer_idx)
        bsz = len(self._batch_rows)
        cum = self.cumulative.view(self.num_layers, self.max_rows, self.width)[kv_idx]
        ops.h2o_normalize_accumulate(self.step_scores[kv_idx, :bsz], 1.0 / math.sqrt(self.head_dim),
                                     cumulative=cum, kv_lens=self._kv_lens_dev[:bsz], width=self.decode_width,
                                     cum_rows=self._rows_dev_i32[:bsz],
                                     refill=SCORE_FILL if self.refill_in_kernel else 0.0)

    # ------------------------------------------------------------------ eviction
    def recent_count(self, kv_len: int) -> int:
        return min(max(1, int(self.budget * self.recent_ratio)), self.budget, kv_len)  # h2o.py:498-499

    def evict_after_decode(self, keep_out: torch.Tensor | None = None) -> bool:
        """Burst eviction (h2o.py:1498-1625): every row of the decode batch that reached budget + interval
        tokens — or, when the step consumed the last free slot, every ACTIVE decode row over budget, scheduled
        or not (1510-1522) — keeps its heavy hitters plus the recent suffix in every layer; slot rows and
        cumulative scores are compacted in place, dropped slots go back to the layer's free stack.  Rows may
        differ in length (the reference groups them by length; here one launch takes per-entry lengths).
        ``keep_out`` [L * evicted rows, budget] receives the kept indices (row order = ``last_evicted_rows``).
        Returns True if an eviction ran."""
        rows = list(self._batch_rows)
        if not rows:
            return False
        self._active_decode_rows.update(rows)
        under_pressure = self.num_free <= 0
        trigger = self.budget + 1 if under_pressure else self.budget + self.interval
        candidates = rows + (sorted(self._active_decode_rows.difference(rows)) if under_pressure else [])
        evict = [r for r in candidates if int(self.row_seq_lens[r]) >= trigger]
        self.last_evicted_rows = evict
        if not evict:
            return False
        self._compact_rows(evict, self.budget, keep_out)
        return True
