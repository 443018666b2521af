"""Quest cache manager of the B200 provider.

Same role and hooks as the reference's ``QuestCacheManager`` (engine/cache_manager/quest.py): paged
KV ``[2, L, slots, Hkv, D]``, a page table per request row, per-page key bounds kept ON THE GPU, and
the decode-time view builder that scores pages against the query, keeps the top pages plus the newest
one and hands attention a packed token-slot table.  What differs is the layout and the execution:

* bounds live in *chunks* of 16 consecutive LOGICAL pages of one row (``ops.bounds_shape``,
  csrc/bounds_layout.cuh) instead of two planes ``[2, L, P, Hkv, D]`` indexed by physical page
  (quest.py:96-105): the scoring scan of a row is a stream of contiguous 8-64 KB reads with no page-table
  indirection and no gathered copy (quest.py:1870-1881 materialises two).  The manager assigns a chunk to
  (row, logical page // 16) when it allocates the page and keeps ``unit_table`` / ``page_rec`` on the device;
* score + mask + top-k + last page + slot packing is ONE launch for long-text batches
  (quest.py:1773-1802, 1883-1913 are ~20 torch launches); mixed batches use two launches;
* a decode token that completes a page refreshes its bounds inside the KV store launch (or, in
  ``decode_bounds_refresh="forward_end"`` mode, one all-layer launch in ``on_forward_end``) instead of
  the per-layer Python loop of quest.py:1719-1771.

Host-side page allocation stays in numpy like the reference (quest.py:1280-1360).
"""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np
import torch

from .. import ops
from ..configs import quest_page_budget
from ..operators.decode_attention import (B200SparseDecodeProvider, SparseDecodeAttentionSpec,
                                          prepare_sparse_decode_attention_op)
from ..utils_context import get_context
from .base import CacheManager, SparseSelection

PAGE = 16


class QuestCacheManager(CacheManager):
    sparse_method = "quest"

    def __init__(self, config, *, num_layers: int, num_kv_heads: int, head_dim: int = 128, num_pages: int,
                 max_rows: int, max_model_len: int, device, dtype: torch.dtype = torch.bfloat16,
                 decode_bounds_refresh: str = "fused", provider: B200SparseDecodeProvider | None = None,
                 fuse_decode_store: bool = True) -> None:
        self.config = config
        self.page_size = int(getattr(config, "quest_chunk_size", PAGE))
        if self.page_size != PAGE:
            raise ValueError(f"the B200 provider supports quest_chunk_size={PAGE} only, got {self.page_size}.")
        if decode_bounds_refresh not in ("fused", "forward_end"):
            raise ValueError("decode_bounds_refresh must be 'fused' or 'forward_end'.")
        self.num_pages = int(num_pages)
        super().__init__(num_layers=num_layers, num_kv_heads=num_kv_heads, head_dim=head_dim,
                         num_slots=self.num_pages * PAGE, device=device, dtype=dtype)
        self.decode_bounds_refresh = decode_bounds_refresh
        self.max_model_len = int(max_model_len)
        self.max_pages_per_row = (self.max_model_len + PAGE - 1) // PAGE
        self.max_rows = int(max_rows)
        # page bounds, resident on the GPU, in chunks of 16 logical pages of one row (a row wastes < 1 chunk)
        self.max_units_per_row = (self.max_pages_per_row + PAGE - 1) // PAGE
        self.num_chunks = (self.num_pages + PAGE - 1) // PAGE + self.max_rows
        self.bounds = self._alloc_bounds()
        self.unit_table = torch.full((self.max_rows, self.max_units_per_row), -1, dtype=torch.int32, device=self.device)
        self.unit_table_cpu = np.full((self.max_rows, self.max_units_per_row), -1, dtype=np.int32)
        self.page_rec = torch.full((self.num_pages,), -1, dtype=torch.int32, device=self.device)
        self.free_chunks_cpu_stack = np.arange(self.num_chunks - 1, -1, -1, dtype=np.int32)
        self._num_free_chunks = self.num_chunks
        self.buffer_req_to_page_slots = torch.full((self.max_rows, self.max_pages_per_row), -1, dtype=torch.int32,
                                                   device=self.device)
        self.buffer_req_to_page_slots_cpu = np.full((self.max_rows, self.max_pages_per_row), -1, dtype=np.int32)
        self.buffer_req_to_token_slots = torch.zeros((self.max_rows, self.max_model_len), dtype=torch.int32,
                                                     device=self.device)
        self.row_seq_lens = np.zeros((self.max_rows,), dtype=np.int32)
        self.seq_id_to_row: dict[int, int] = {}
        self._free_rows = list(range(self.max_rows - 1, -1, -1))
        self.free_pages_cpu_stack = np.arange(self.num_pages, dtype=np.int32)
        self._num_free_pages = self.num_pages
        self._decode_static_max_context_len: int | None = None
        self._scratch: dict[int, SimpleNamespace] = {}
        self._provider = provider
        # decode: the layer's new K/V rows wait here for the layer's scoring launch, which stores them itself
        self.fuse_decode_store = bool(fuse_decode_store)
        # Sparse long-text layers as ONE launch (decode_layer_one_launch): True / False / "auto".  Measured on B200
        # (profiles/r2_d_evict_first.txt): the one-launch layer wins where a CTA streams >= ~1.5 MB of bounds (cfg5: 147.6
        # vs 152.3 us per layer, cfg3: 143.6 vs 154.4, the N=2 shard of cfg5: 91.7 vs 95.9); it loses on small launches
        # (N=4 / N=8 shards: 66.4 vs 64.1, 54.0 vs 51.4; cfg2: 40.6
        # vs 39.5), where the selection tail and the flag -> slots -> data chain are not amortised.
        self.fuse_decode_layer = "auto"
        self._pending_store: tuple[int, ops.DecodeStore] | None = None

    def _alloc_bounds(self) -> torch.Tensor:
        if self.device.type != "cuda":  # host-logic tests (_allow_host_tensors): same shape, no kernel touches it
            return torch.zeros((self.num_layers, *ops.bounds_shape(self.num_chunks, self.num_kv_heads)), dtype=torch.bfloat16)
        return ops.alloc_bounds(self.num_chunks, self.num_kv_heads, self.device, layers=self.num_layers)

    def bounds_by_page(self, layer_idx: int) -> torch.Tensor:
        """``[num_pages, 2 (max, min), Hkv, D]`` indexed by physical page — the reference's ``metadata_cache``
        order (quest.py:96-105) — rebuilt from the chunk layout (debugging / tests, not the hot path)."""
        return ops.unpack_bounds(self.bounds[self.kv_layer_index(layer_idx)], self.page_rec)

    # ------------------------------------------------------------------ provider / scratch
    def _get_provider(self, num_heads: int) -> B200SparseDecodeProvider:
        if self._provider is None:
            spec = SparseDecodeAttentionSpec(int(num_heads), self.num_kv_heads, self.head_dim, torch.bfloat16,
                                             page_size=PAGE, sparse_method="quest")
            self._provider = prepare_sparse_decode_attention_op(spec, device_index=self.device.index).provider
        return self._provider

    def _get_scratch(self, batch: int, n_cols: int, width: int) -> SimpleNamespace:
        key = int(batch)
        s = self._scratch.get(key)
        stride = ops.score_row_stride(n_cols)
        if s is None or s.scores.shape[1] < stride or s.packed.shape[1] < width:
            s = SimpleNamespace(
                scores=torch.empty((batch, stride), dtype=torch.bfloat16, device=self.device),
                packed=torch.zeros((batch, width), dtype=torch.int32, device=self.device),
                lens=torch.zeros((batch,), dtype=torch.int32, device=self.device),
                tickets=ops.sched_scratch(batch, self.device),
                arange=torch.arange(batch, dtype=torch.int32, device=self.device))
            self._scratch[key] = s
        return s

    # ------------------------------------------------------------------ allocation (host)
    @property
    def num_free_slots(self) -> int:
        return self._num_free_pages * PAGE

    def _get_free_row(self, seq_id: int) -> int:
        row = self.seq_id_to_row.get(seq_id)
        if row is None:
            if not self._free_rows:
                raise RuntimeError("Out of QuEST request rows.")
            row = self._free_rows.pop()
            self.seq_id_to_row[seq_id] = row
            self.row_seq_lens[row] = 0
        return row

    def _h2d(self, arr, dtype: torch.dtype = torch.int64) -> torch.Tensor:
        """Small host array -> device tensor WITHOUT synchronising the host with the stream: staged through a ring of
        pinned buffers and copied non-blocking (a `.to(device)` of pageable memory, or a blocking `copy_`, waits for
        everything queued on the stream, i.e. for the whole previous decode step: the host and the GPU then take
        turns).  A staging buffer is reused only after the copy out of it has run (event per buffer)."""
        a = np.ascontiguousarray(arr)
        if self.device.type != "cuda":
            return torch.from_numpy(a).to(dtype).to(self.device)
        ring = getattr(self, "_stage_ring", None)
        if ring is None:
            ring = self._stage_ring = SimpleNamespace(
                bufs=[torch.empty(8192, dtype=torch.int64).pin_memory() for _ in range(64)],
                events=[None] * 64, nxt=0)
        n = int(a.size)
        if n > ring.bufs[0].numel():
            return torch.from_numpy(a).to(dtype).to(self.device)  # (not a decode-step size)
        i = ring.nxt
        ring.nxt = (i + 1) % len(ring.bufs)
        if ring.events[i] is not None:
            ring.events[i].synchronize()
        view = ring.bufs[i].view(dtype)[:n]
        view.numpy()[:] = a.reshape(-1)
        out = torch.empty(n, dtype=dtype, device=self.device)
        out.copy_(view, non_blocking=True)
        ev = ring.events[i] or torch.cuda.Event()
        ev.record()
        ring.events[i] = ev
        return out.view(a.shape)

    def _pop_pages(self, n: int) -> np.ndarray:
        if self._num_free_pages < n:
            raise RuntimeError(f"Out of QuEST KV pages: need_pages={n}, free_pages={self._num_free_pages}.")
        ptr = self._num_free_pages
        pages = self.free_pages_cpu_stack[ptr - n: ptr][::-1].copy()
        self._num_free_pages -= n
        return pages

    def _assign_bounds_records(self, rows: np.ndarray, page_idx: np.ndarray, pages: np.ndarray) -> None:
        """Give the freshly allocated ``pages`` (logical page ``page_idx`` of table row ``rows``) their bounds
        records: a new chunk per (row, page_idx // 16) when page_idx % 16 == 0.  Host numpy + three small
        H2D index updates, like the page-table update next to it."""
        rows = np.asarray(rows, dtype=np.int64)
        page_idx = np.asarray(page_idx, dtype=np.int64)
        units = page_idx // PAGE
        new_unit = np.nonzero(page_idx % PAGE == 0)[0]
        if new_unit.size:
            if self._num_free_chunks < new_unit.size:
                raise RuntimeError(f"Out of QuEST bounds chunks: need={new_unit.size}, free={self._num_free_chunks}.")
            ptr = self._num_free_chunks
            chunks = self.free_chunks_cpu_stack[ptr - new_unit.size: ptr][::-1].copy()
            self._num_free_chunks -= int(new_unit.size)
            self.unit_table_cpu[rows[new_unit], units[new_unit]] = chunks
            self.unit_table[self._h2d(rows[new_unit]), self._h2d(units[new_unit])] = \
                self._h2d(chunks.astype(np.int32), torch.int32)
        recs = self.unit_table_cpu[rows, units].astype(np.int64) * PAGE + page_idx % PAGE
        self.page_rec[self._h2d(np.asarray(pages, dtype=np.int64))] = self._h2d(recs.astype(np.int32), torch.int32)

    def allocate_prefill(self, seq_id: int, num_tokens: int) -> torch.Tensor:
        """Slots for the next ``num_tokens`` tokens of a sequence (chunked prefill, quest.py:1426-1493)."""
        row = self._get_free_row(seq_id)
        cur = int(self.row_seq_lens[row])
        new_len = cur + int(num_tokens)
        if new_len > self.max_model_len:
            raise RuntimeError(f"KV row length exceeds max_model_len: {new_len} > {self.max_model_len}")
        have = (cur + PAGE - 1) // PAGE
        need = (new_len + PAGE - 1) // PAGE
        if need > have:
            pages = self._pop_pages(need - have)
            self.buffer_req_to_page_slots_cpu[row, have:need] = pages
            self.buffer_req_to_page_slots[row, have:need] = torch.from_numpy(pages).to(self.device)
            self._assign_bounds_records(np.full(need - have, row), np.arange(have, need), pages)
        pos = np.arange(cur, new_len, dtype=np.int64)
        slots = self.buffer_req_to_page_slots_cpu[row, pos // PAGE].astype(np.int64) * PAGE + pos % PAGE
        slots_t = torch.from_numpy(slots.astype(np.int32)).to(self.device)
        self.buffer_req_to_token_slots[row, cur:new_len] = slots_t
        self.row_seq_lens[row] = new_len
        self.layer_batch_state.slot_mapping = slots_t
        return slots_t

    def free_seq(self, seq_id: int) -> None:
        row = self.seq_id_to_row.pop(seq_id, None)
        if row is None:
            return
        n_pages = (int(self.row_seq_lens[row]) + PAGE - 1) // PAGE
        pages = self.buffer_req_to_page_slots_cpu[row, :n_pages]
        ptr = self._num_free_pages
        self.free_pages_cpu_stack[ptr: ptr + n_pages] = pages
        self._num_free_pages += n_pages
        self.page_rec[torch.from_numpy(pages.astype(np.int64)).to(self.device)] = -1
        n_units = (n_pages + PAGE - 1) // PAGE
        ptr = self._num_free_chunks
        self.free_chunks_cpu_stack[ptr: ptr + n_units] = self.unit_table_cpu[row, :n_units]
        self._num_free_chunks += n_units
        self.unit_table_cpu[row, :] = -1
        self.unit_table[row, :] = -1
        self.buffer_req_to_page_slots_cpu[row, :] = -1
        self.buffer_req_to_page_slots[row, :] = -1
        self.row_seq_lens[row] = 0
        self._free_rows.append(row)

    def _allocate_batch(self, seq_ids: list[int]) -> tuple[np.ndarray, np.ndarray]:
        """One slot per sequence (decode); a new page is popped when len % 16 == 0 (quest.py:1280-1360)."""
        rows = np.asarray([self._get_free_row(s) for s in seq_ids], dtype=np.int64)
        cur = self.row_seq_lens[rows].astype(np.int64)
        if cur.size and int(cur.max()) + 1 > self.max_model_len:
            raise RuntimeError(f"KV row length exceeds max_model_len in QuEST _allocate_batch: max_cur_len={int(cur.max())}")
        page_idx, page_off = cur // PAGE, cur % PAGE
        new_pos = np.nonzero(page_off == 0)[0]
        if new_pos.size:
            pages = self._pop_pages(int(new_pos.size))
            self.buffer_req_to_page_slots_cpu[rows[new_pos], page_idx[new_pos]] = pages
            self.buffer_req_to_page_slots[self._h2d(rows[new_pos]), self._h2d(page_idx[new_pos])] = \
                self._h2d(np.asarray(pages, dtype=np.int32), torch.int32)
            self._assign_bounds_records(rows[new_pos], page_idx[new_pos], pages)
        slots = self.buffer_req_to_page_slots_cpu[rows, page_idx].astype(np.int64) * PAGE + page_off
        slots32 = slots.astype(np.int32)
        self.buffer_req_to_token_slots[self._h2d(rows), self._h2d(cur)] = self._h2d(slots32, torch.int32)
        self.row_seq_lens[rows] += 1
        return rows, slots32

    def set_decode_static_max_context_len(self, cap: int | None) -> None:
        """Pin the context bucket of the static / graph decode runner (decode_cuda_graph.py:16-26,282-300)."""
        self._decode_static_max_context_len = None if cap is None else int(cap)

    def prepare_decode_static(self, seqs, input_ids, positions, slot_mapping, context_lens, req_indices):
        """Fill caller-owned static buffers for one decode step (quest.py:1543-1605).  Padding entries
        replicate row 0 with slot -1."""
        real = len(seqs)
        graph_bs = int(input_ids.numel())
        if real <= 0:
            raise ValueError("Static decode requires a non-empty real decode batch.")
        if any(int(t.numel()) != graph_bs for t in (positions, slot_mapping, context_lens, req_indices)):
            raise ValueError("Static decode metadata buffers must have the same graph batch size.")
        if real > graph_bs:
            raise ValueError(f"Static decode graph batch is smaller than the real decode batch: graph={graph_bs}, real={real}.")
        seq_ids = [s.seq_id for s in seqs]
        rows, slots = self._allocate_batch(seq_ids)
        lens = self.row_seq_lens[rows].astype(np.int32)
        host = np.empty((3, graph_bs), dtype=np.int32)
        host[0, :real], host[1, :real], host[2, :real] = slots, lens, rows
        host[0, real:], host[1, real:], host[2, real:] = -1, lens[0], rows[0]
        ids = [int(s.decode_input_token) for s in seqs] + [int(seqs[0].decode_input_token)] * (graph_bs - real)
        pos = [int(s.decode_input_position) for s in seqs] + [int(seqs[0].decode_input_position)] * (graph_bs - real)
        if slot_mapping.device.type == "cuda":
            # one staged, non-blocking copy for the five buffers (see _h2d: nothing here waits for the stream)
            meta = np.empty((5, graph_bs), dtype=np.int64)
            meta[:3], meta[3], meta[4] = host, ids, pos
            dev = self._h2d(meta)
            slot_mapping.copy_(dev[0])
            context_lens.copy_(dev[1])
            req_indices.copy_(dev[2])
            input_ids.copy_(dev[3])
            positions.copy_(dev[4])
        else:
            packed = torch.from_numpy(host)
            slot_mapping.copy_(packed[0])
            context_lens.copy_(packed[1])
            req_indices.copy_(packed[2])
            input_ids.copy_(torch.tensor(ids, dtype=input_ids.dtype))
            positions.copy_(torch.tensor(pos, dtype=positions.dtype))
        st = self.layer_batch_state
        st.slot_mapping, st.context_lens, st.req_indices = slot_mapping, context_lens, req_indices
        st.max_context_len = int(lens.max())
        if self._decode_static_max_context_len is not None:
            st.max_context_len = max(st.max_context_len, self._decode_static_max_context_len)
        return input_ids, positions, None

    # ------------------------------------------------------------------ KV store
    def save_rope_kv_if_needed(self, layer_idx: int, k: torch.Tensor, v: torch.Tensor,
                               slot_mapping: torch.Tensor | None = None) -> None:
        """Decode step of a sparse layer with a long-text batch: the rows are handed to the layer's scoring
        launch (``build_decode_view`` follows in the same layer; a decode token lands in its row's last page,
        which that launch never scores) instead of a launch of their own.  Anything that looks at the cache
        before (``get_layer_kv_cache``, the next store, ``on_forward_end``) stores them first."""
        self.flush_pending_store()
        if slot_mapping is None:
            slot_mapping = self.layer_batch_state.slot_mapping
        ctx = get_context()
        if (self.fuse_decode_store and not ctx.is_prefill and bool(ctx.is_long_text)
                and layer_idx >= int(self.config.quest_skip_layers) and int(self.config.quest_token_budget) > 0
                and slot_mapping is not None and int(k.shape[0]) == int(slot_mapping.numel())):
            k_cache, v_cache = CacheManager.get_layer_kv_cache(self, layer_idx)
            self._pending_store = (layer_idx, ops.DecodeStore(k, v, k_cache, v_cache, slot_mapping,
                                                              self.decode_bounds_refresh == "fused", self.page_rec))
            return
        super().save_rope_kv_if_needed(layer_idx, k, v, slot_mapping)

    def flush_pending_store(self) -> None:
        pend, self._pending_store = self._pending_store, None
        if pend is not None:
            layer_idx, st = pend
            ops.store_kv_minmax(st.key, st.value, st.k_cache, st.v_cache, st.slot_mapping,
                                self.bounds[self.kv_layer_index(layer_idx)] if st.refresh_bounds else None, self.page_rec)

    def _take_pending_store(self, layer_idx: int) -> ops.DecodeStore | None:
        if self._pending_store is None:
            return None
        if self._pending_store[0] != layer_idx:
            self.flush_pending_store()
            return None
        st = self._pending_store[1]
        self._pending_store = None
        return st

    def get_layer_kv_cache(self, layer_idx: int) -> tuple[torch.Tensor, torch.Tensor]:
        self.flush_pending_store()
        return super().get_layer_kv_cache(layer_idx)

    # ------------------------------------------------------------------ bounds maintenance
    def _fused_store_bounds(self, layer_idx: int):
        if get_context().is_prefill or self.decode_bounds_refresh != "fused":
            return None
        return self.bounds[self.kv_layer_index(layer_idx)]

    def on_kv_stored(self, layer_idx: int, k: torch.Tensor, slot_mapping: torch.Tensor) -> None:
        """Prefill: bounds of every page completed by this chunk (quest.py:1608-1716).  Decode: nothing here
        (the store launch or on_forward_end refreshes a completed page)."""
        if slot_mapping is None or slot_mapping.numel() == 0 or not get_context().is_prefill:
            return
        kv_idx = self.kv_layer_index(layer_idx)
        ops.quest_page_minmax(self.kv_cache[0, kv_idx], self.bounds[kv_idx], self.page_rec, slot_mapping)

    def on_forward_end(self, seqs, is_prefill: bool) -> None:
        """quest.py:1719-1771: after a decode step, rebuild the bounds of pages it completed (all layers)."""
        self.flush_pending_store()
        if is_prefill or self.decode_bounds_refresh != "forward_end":
            return
        st = self.layer_batch_state
        ops.quest_decode_page_complete(self.kv_cache[0], self.bounds, self.page_rec, self.buffer_req_to_page_slots,
                                       st.req_indices, st.context_lens)

    # ------------------------------------------------------------------ decode view
    def _default_active_slots_for_selection(self, layer_idx: int, selection: SparseSelection) -> torch.Tensor:
        return self.buffer_req_to_token_slots

    def build_decode_view(self, layer_idx, q, active_slots, req_indices, context_lens, *, num_heads, num_kv_heads):
        """quest.py:1805-1830."""
        if layer_idx < int(self.config.quest_skip_layers):
            return active_slots, req_indices, context_lens
        token_budget = int(self.config.quest_token_budget)
        if token_budget <= 0:
            return active_slots, req_indices, context_lens
        return self._build_decode_view_static(layer_idx, q, active_slots, req_indices, context_lens,
                                              token_budget=token_budget, num_heads=num_heads,
                                              num_kv_heads=num_kv_heads)

    @torch.no_grad()
    def decode_layer_one_launch(self, layer_idx, q, selection, *, num_heads, num_kv_heads, out, workspace,
                                tp_comm=None, tp_gathered=None):
        """A sparse layer of a long-text decode batch in ONE launch (svb200_quest_decode_layer): the pending K/V store,
        `_build_decode_view_static` and the decode attention.  Returns the attention output, or None when the layer
        takes the general path (dense / skip layer, mixed batch, group size > 8, `fuse_decode_layer` off)."""
        if not self.fuse_decode_layer or layer_idx < int(self.config.quest_skip_layers):
            return None
        if self.fuse_decode_layer == "auto":
            bounds_bytes = int(q.shape[0]) * self.max_pages_per_row * 2 * self.num_kv_heads * 128 * 2
            if bounds_bytes < 148 * (3 << 19):
                return None
        token_budget = int(self.config.quest_token_budget)
        ctx = get_context()
        if token_budget <= 0 or ctx.is_prefill or not bool(ctx.is_long_text) or int(num_heads) // int(num_kv_heads) > 8:
            return None
        if selection.kind != "full" or int(num_kv_heads) != self.num_kv_heads:
            return None
        page_budget_base = max(3, token_budget // PAGE)
        max_keep = max(token_budget, page_budget_base * PAGE, PAGE)
        max_context_len = self.layer_batch_state.max_context_len
        if max_context_len is None or int(max_context_len) <= max_keep:
            return None
        max_pages = min(self.max_pages_per_row, (int(max_context_len) + PAGE - 1) // PAGE)
        _, prev_budget = quest_page_budget(token_budget, PAGE, max_pages)
        if prev_budget <= 0 or max_pages < 2:
            return None
        batch = int(q.shape[0])
        n_cols = max_pages - 1
        s = self._get_scratch(batch, n_cols, (prev_budget + 1) * PAGE)
        store = self._take_pending_store(layer_idx)
        kv = self.kv_layer_index(layer_idx)
        if out is None:
            out = torch.empty_like(q)
        return self._get_provider(num_heads).quest_decode_layer(
            q, self.bounds[kv], self.unit_table, self.buffer_req_to_page_slots, selection.req_indices,
            selection.context_lens, n_cols=n_cols, num_kv_heads=self.num_kv_heads, prev_budget=prev_budget,
            scores=s.scores, packed_slots=s.packed, sparse_lens=s.lens, tickets=s.tickets,
            k_cache=self.kv_cache[0, kv], v_cache=self.kv_cache[1, kv], out=out, workspace=workspace, store=store,
            comm=tp_comm, gathered=tp_gathered)

    @torch.no_grad()
    def _build_decode_view_static(self, layer_idx, q, active_slots, req_indices, context_lens, *, token_budget,
                                  num_heads, num_kv_heads):
        """quest.py:1833-1913 in one (long-text) or two (mixed) launches."""
        page_budget_base = max(3, int(token_budget) // PAGE)
        max_keep = max(int(token_budget), page_budget_base * PAGE, PAGE)
        max_context_len = self.layer_batch_state.max_context_len
        if max_context_len is None:
            raise RuntimeError("QuEST decode CUDA graph requires max_context_len to be pinned.")
        max_context_len = int(max_context_len)
        if max_context_len <= max_keep:
            return active_slots, req_indices, context_lens
        batch = int(q.shape[0])
        max_pages = min(self.max_pages_per_row, (max_context_len + PAGE - 1) // PAGE)
        _, prev_budget = quest_page_budget(token_budget, PAGE, max_pages)
        if prev_budget <= 0:
            return active_slots, req_indices, context_lens
        if int(num_kv_heads) != self.num_kv_heads:
            raise ValueError(f"num_kv_heads={num_kv_heads} does not match the cache ({self.num_kv_heads}).")
        n_cols = max_pages - 1
        long_text = bool(get_context().is_long_text)
        width = (prev_budget + 1) * PAGE if long_text else max_keep
        s = self._get_scratch(batch, n_cols, width)
        self._get_provider(num_heads).quest_select(
            q, self.bounds[self.kv_layer_index(layer_idx)], self.unit_table, self.buffer_req_to_page_slots, req_indices,
            context_lens, n_cols=n_cols,
            num_kv_heads=self.num_kv_heads, prev_budget=prev_budget, scores=s.scores, packed_slots=s.packed,
            sparse_lens=s.lens, tickets=s.tickets,
            dense_token_table=None if long_text else self.buffer_req_to_token_slots,
            token_budget=int(token_budget), page_budget_base=page_budget_base, out_width=width,
            store=self._take_pending_store(layer_idx))
        return s.packed[:, :width], s.arange, s.lens
