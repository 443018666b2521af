"""Reference-side binding of libsvb200.so: the file a Sparse-vLLM maintainer would add as
``sparsevllm/kernels/external/svb200.py``.  It imports the REFERENCE package (`sparsevllm`) and
subclasses its own classes; it is not part of the provider package and nothing in sparse-vllm_b200/ imports it.

* ``SVB200AttentionBackend(TritonAttentionBackend)``: ``run_decode`` (layers/attention_backend.py:217-349) ->
  one ``svb200_sparse_decode_attn`` launch (stage 1 (+2-D score) and stage 2 of the Triton path).
* ``QuestB200CacheManager(QuestCacheManager)``: bounds kept in the provider's chunk layout next to the reference's
  page allocator (``_allocate`` 1228-1277, ``_allocate_batch`` 1280-1360, ``free_seq`` 1379-1420 give every new
  page its bounds record), ``on_kv_stored`` (1608-1716) -> ``svb200_quest_page_minmax``, ``on_forward_end``
  (1719-1771) -> ``svb200_quest_decode_page_complete``, ``_build_decode_view_static`` (1833-1913) ->
  ``svb200_quest_score_select_pack``.  ``layers/attention.py`` and ``method_registry.py`` stay untouched.

tests/test_integration_reference.py executes these classes against the real reference on CPU tensors with the
library calls replaced by a recorder that checks every call against the C prototypes.
"""
from __future__ import annotations

import ctypes
from types import SimpleNamespace

import numpy as np
import torch

from sparsevllm.engine.cache_manager.quest import QuestCacheManager
from sparsevllm.layers.attention_backend import TritonAttentionBackend
from sparsevllm.utils.context import get_context
from sparsevllm.utils.profiler import profiler

PAGE = 16
HEAD_DIM = 128
ABI_VERSION = 6
P, I32, I64, F32 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64, ctypes.c_float
# argument types of the entry points bound here (the complete table is sparse-vllm_b200/_native.py:_SIGNATURES)
ARGTYPES = {
    "svb200_workspace_bytes": [I32, I32, I32],
    "svb200_sched_scratch_ints": [I32],
    "svb200_quest_page_minmax": [P, P, I64, P, P, I32, I32, I64, P],
    "svb200_quest_decode_page_complete": [P, P, I64, P, P, I64, P, P, I32, I32, I32, I64, I64, P],
    "svb200_quest_page_score": [P, I64, P, I64, P, I64, P, P, P, I64, I32, I32, I32, I32, P, P],
    "svb200_quest_select_pack": [P, I64, I32, P, I64, P, P, I32, I32, P, I64, P, P, P, I64, I32, I32, I32, P],
    "svb200_quest_score_select_pack": [P, I64, P, I64, P, I64, P, I64, P, P, P, I64, I32, I32, I32, I32, I32, P, I64,
                                       P, P, P, P],
    "svb200_sparse_decode_attn": [P, I64, P, P, I64, P, I64, I32, P, P, P, I64, P, I64, I32, I32, I32, I32, P, I64,
                                  F32, P],
}
_lib = None


def load(path: str = "libsvb200.so"):
    """dlopen the provider library and declare the prototypes used below."""
    global _lib
    lib = ctypes.CDLL(path)
    if lib.svb200_abi_version() != ABI_VERSION:
        raise RuntimeError(f"libsvb200.so ABI {lib.svb200_abi_version()} != {ABI_VERSION}")
    lib.svb200_last_error.restype = ctypes.c_char_p
    for name, argtypes in ARGTYPES.items():
        fn = getattr(lib, name)
        fn.argtypes = argtypes
        fn.restype = I64 if name in ("svb200_workspace_bytes", "svb200_sched_scratch_ints") else I32
    _lib = lib
    return lib


def _check(rc: int, what: str) -> None:
    if rc != 0:  # the reference is fail-fast (AGENTS.md:41-73): no fallback to the Triton path
        raise RuntimeError(f"{what} failed ({rc}): {_lib.svb200_last_error().decode()}")


def _stream(t: torch.Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream if t.is_cuda else None


def bounds_head_split(hkv: int) -> int:
    for d in range(1, hkv):
        if hkv % d == 0 and hkv // d <= 4:
            return d
    return 1 if hkv <= 4 else hkv


class SVB200AttentionBackend(TritonAttentionBackend):
    """Second backend next to TritonAttentionBackend; ``Attention.forward`` is untouched."""

    _workspaces: dict = {}

    def _workspace(self, batch: int, num_heads: int, num_kv_heads: int, device) -> torch.Tensor:
        key = (batch, num_heads, num_kv_heads, str(device))
        ws = self._workspaces.get(key)
        if ws is None:  # zeroed once per shape, outside graph capture; launches leave the ticket part zeroed
            nbytes = int(_lib.svb200_workspace_bytes(batch, num_heads, num_kv_heads))
            ws = torch.zeros((nbytes + 3) // 4, dtype=torch.int32, device=device)
            self._workspaces[key] = ws
        return ws

    def run_decode(self, q, view, *, mid_o, mid_o_logexpsum, max_len_in_batch, block_seq, num_heads, num_kv_heads,
                   gqa_block_n=16, gqa_num_warps=2):
        del mid_o, mid_o_logexpsum, block_seq, gqa_block_n, gqa_num_warps  # the Triton split-K workspace / knobs
        meta, payload = view.meta, view.payload  # AttentionViewMeta / ExplicitKVPayload (cache_manager/base.py:129-211)
        ws = self._workspace(int(q.shape[0]), int(num_heads), int(num_kv_heads), q.device)
        o = torch.empty_like(q)
        score = meta.attn_score  # None (Quest) or fp32 [B, W] pre-filled with -1e20 (H2O)
        # same profiler tag family as attention_backend.py:284-347 (stage 1 and 2 are one launch here)
        kind = "full" if int(max_len_in_batch) > 8192 else "sparse"  # attention_backend.py:284
        with profiler.record(f"decode_attention_stage1_{kind}"):
            self._launch(q, meta, payload, o, score, ws, max_len_in_batch, num_heads, num_kv_heads)
        return o

    @staticmethod
    def _launch(q, meta, payload, o, score, ws, max_len_in_batch, num_heads, num_kv_heads):
        _check(_lib.svb200_sparse_decode_attn(
            q.data_ptr(), q.stride(0), payload.k_cache.data_ptr(), payload.v_cache.data_ptr(),
            int(payload.k_cache.shape[0]), meta.active_slots.data_ptr(), meta.active_slots.stride(0),
            int(max_len_in_batch), meta.req_indices.data_ptr(), meta.context_lens.data_ptr(),
            o.data_ptr(), o.stride(0), score.data_ptr() if score is not None else None,
            score.stride(0) if score is not None else 0, int(q.shape[0]), int(num_heads), int(num_kv_heads), 0,
            ws.data_ptr(), ws.numel() * 4, float(q.shape[-1]) ** -0.5, _stream(q)), "svb200_sparse_decode_attn")


class QuestB200CacheManager(QuestCacheManager):
    """Registered under the same ``sparse_method`` name; ``CacheManager.create`` (base.py:329-369) returns it when
    the resolved decode provider is ``b200_sm100a_sparse_decode``."""

    # ---- state -------------------------------------------------------------------------------------------
    def allocate_kv_cache(self):
        super().allocate_kv_cache()
        self.metadata_cache = None  # replaced by the chunk layout below
        self._init_b200_state()

    def _init_b200_state(self) -> None:
        hkv, dev = int(self.num_kv_heads), self.device
        hs = bounds_head_split(hkv)
        rows = int(self.buffer_req_to_page_slots.shape[0])
        self.max_units_per_row = (int(self.max_pages_per_row) + PAGE - 1) // PAGE
        self.num_chunks = (int(self.num_pages) + PAGE - 1) // PAGE + rows
        self.bounds = torch.zeros((int(self.num_kv_layers), self.num_chunks, hs, PAGE, 2, hkv // hs, HEAD_DIM),
                                  dtype=torch.bfloat16, device=dev)
        self.unit_table = torch.full((rows, self.max_units_per_row), -1, dtype=torch.int32, device=dev)
        self.unit_table_cpu = np.full((rows, self.max_units_per_row), -1, dtype=np.int32)
        self.page_rec = torch.full((int(self.num_pages),), -1, dtype=torch.int32, device=dev)
        self._free_chunks = list(range(self.num_chunks - 1, -1, -1))
        self._row_pages_mapped = np.zeros((rows,), dtype=np.int64)
        self._select_scratch: dict[int, SimpleNamespace] = {}

    # ---- page allocation: every new page gets its bounds record --------------------------------------------
    def _sync_bounds_records(self, rows) -> None:
        for row in np.unique(np.asarray(rows, dtype=np.int64)):
            have = int(self._row_pages_mapped[row])
            need = (int(self.row_seq_lens[row]) + PAGE - 1) // PAGE
            if need <= have:
                continue
            idx = np.arange(have, need)
            for u in np.unique(idx // PAGE):
                if self.unit_table_cpu[row, u] < 0:
                    self.unit_table_cpu[row, u] = self._free_chunks.pop()
                    self.unit_table[row, int(u)] = int(self.unit_table_cpu[row, u])
            pages = self.buffer_req_to_page_slots_cpu[row, have:need].astype(np.int64)
            recs = self.unit_table_cpu[row, idx // PAGE].astype(np.int64) * PAGE + idx % PAGE
            self.page_rec[torch.from_numpy(pages).to(self.device)] = torch.from_numpy(recs.astype(np.int32)).to(self.device)
            self._row_pages_mapped[row] = need

    def _allocate(self, seq_id: int, size: int):
        slots = super()._allocate(seq_id, size)
        self._sync_bounds_records([self.seq_id_to_row[seq_id]])
        return slots

    def _allocate_batch(self, seq_ids, size, *, graph_batch_size=None):
        slots = super()._allocate_batch(seq_ids, size, graph_batch_size=graph_batch_size)
        self._sync_bounds_records([self.seq_id_to_row[s] for s in seq_ids])
        return slots

    def free_seq(self, seq_id: int):
        row = self.seq_id_to_row.get(seq_id)
        if row is not None:
            n = int(self._row_pages_mapped[row])
            if n:
                pages = self.buffer_req_to_page_slots_cpu[row, :n].astype(np.int64)
                self.page_rec[torch.from_numpy(pages).to(self.device)] = -1
            used = self.unit_table_cpu[row]
            self._free_chunks.extend(int(c) for c in used[used >= 0])
            self.unit_table_cpu[row, :] = -1
            self.unit_table[row, :] = -1
            self._row_pages_mapped[row] = 0
        return super().free_seq(seq_id)

    # ---- bounds maintenance --------------------------------------------------------------------------------
    def on_kv_stored(self, layer_idx: int, k: torch.Tensor, slot_mapping: torch.Tensor):
        if slot_mapping is None or slot_mapping.numel() == 0:
            return
        kv_idx = self.kv_layer_index(layer_idx)
        k_cache = self.kv_cache[0, kv_idx]
        with profiler.record("quest_update_metadata"):  # quest.py:1628
            self._page_minmax(k_cache, kv_idx, k, slot_mapping)

    def _page_minmax(self, k_cache, kv_idx, k, slot_mapping):
        _check(_lib.svb200_quest_page_minmax(
            k_cache.data_ptr(), self.bounds[kv_idx].data_ptr(), self.num_chunks, self.page_rec.data_ptr(),
            slot_mapping.data_ptr(), int(slot_mapping.numel()), int(self.num_kv_heads), int(k_cache.shape[0]),
            _stream(k)), "svb200_quest_page_minmax")

    def on_forward_end(self, seqs, is_prefill: bool):
        if is_prefill or not seqs:
            return
        st = self.layer_batch_state
        with profiler.record("quest_update_metadata_decode_pages"):  # quest.py:1741
            self._decode_page_complete(st)

    def _decode_page_complete(self, st):
        _check(_lib.svb200_quest_decode_page_complete(
            self.kv_cache[0].data_ptr(), self.bounds.data_ptr(), self.num_chunks, self.page_rec.data_ptr(),
            self.buffer_req_to_page_slots.data_ptr(), self.buffer_req_to_page_slots.stride(0),
            st.req_indices.data_ptr(), st.context_lens.data_ptr(), int(st.req_indices.numel()),
            int(self.num_kv_layers), int(self.num_kv_heads), int(self.kv_cache.shape[2]), int(self.num_pages),
            _stream(self.bounds)), "svb200_quest_decode_page_complete")

    # ---- decode view ---------------------------------------------------------------------------------------
    def _scratch(self, batch: int, n_cols: int, width: int) -> SimpleNamespace:
        s = self._select_scratch.get(batch)
        stride = (n_cols + 7) // 8 * 8
        if s is None or s.scores.shape[1] < stride or s.packed.shape[1] < width:
            dev = self.device
            s = SimpleNamespace(
                scores=torch.empty((batch, stride), dtype=torch.bfloat16, device=dev),
                packed=torch.zeros((batch, width), dtype=torch.int32, device=dev),
                lens=torch.zeros((batch,), dtype=torch.int32, device=dev),
                tickets=torch.zeros((int(_lib.svb200_sched_scratch_ints(batch)),), dtype=torch.int32, device=dev),
                arange=torch.arange(batch, dtype=torch.int32, device=dev))
            self._select_scratch[batch] = s
        return s

    @torch.no_grad()
    def _build_decode_view_static(self, layer_idx, q, active_slots, req_indices, context_lens, *, token_budget,
                                  num_kv_heads):
        with profiler.record("quest_build_decode_view_static"):  # quest.py:1844
            return self._decode_view(layer_idx, q, active_slots, req_indices, context_lens, token_budget, num_kv_heads)

    def _decode_view(self, layer_idx, q, active_slots, req_indices, context_lens, token_budget, num_kv_heads):
        num_heads = int(q.shape[1])
        page = int(self.page_size)
        page_budget_base = max(3, int(token_budget) // page)
        max_keep = max(int(token_budget), page_budget_base * page, page)
        max_context_len = self.layer_batch_state.max_context_len
        if max_context_len is None:
            raise RuntimeError("QuEST decode CUDA graph requires max_context_len to be pinned.")
        if int(max_context_len) <= max_keep:
            return active_slots, req_indices, context_lens
        max_pages = min(int(self.max_pages_per_row), (int(max_context_len) + page - 1) // page)
        prev_budget = min(page_budget_base - 1, max_pages - 1)
        if prev_budget <= 0:
            return active_slots, req_indices, context_lens
        batch, n_cols = int(q.shape[0]), max_pages - 1
        long_text = bool(get_context().is_long_text)
        s = self._scratch(batch, n_cols, (prev_budget + 1) * page if long_text else max_keep)
        kv_idx = self.kv_layer_index(layer_idx)
        pt = self.buffer_req_to_page_slots
        if not long_text:  # mixed short/long rows (quest.py:1866-1868, 1903-1911): score, then select with the dense merge
            _check(_lib.svb200_quest_page_score(
                q.data_ptr(), q.stride(0), self.bounds[kv_idx].data_ptr(), self.num_chunks, self.unit_table.data_ptr(),
                self.unit_table.stride(0), req_indices.data_ptr(), context_lens.data_ptr(), s.scores.data_ptr(),
                s.scores.stride(0), batch, int(num_heads), int(num_kv_heads), n_cols, s.tickets.data_ptr(), _stream(q)),
                "svb200_quest_page_score")
            dense = self.buffer_req_to_token_slots
            _check(_lib.svb200_quest_select_pack(
                s.scores.data_ptr(), s.scores.stride(0), n_cols, pt.data_ptr(), pt.stride(0), req_indices.data_ptr(),
                context_lens.data_ptr(), batch, prev_budget, s.packed.data_ptr(), s.packed.stride(0), s.lens.data_ptr(),
                None, dense.data_ptr(), dense.stride(0), int(token_budget), page_budget_base, max_keep, _stream(q)),
                "svb200_quest_select_pack")
            return s.packed[:, :max_keep], s.arange, s.lens
        _check(_lib.svb200_quest_score_select_pack(
            q.data_ptr(), q.stride(0), self.bounds[kv_idx].data_ptr(), self.num_chunks,
            self.unit_table.data_ptr(), self.unit_table.stride(0),
            self.buffer_req_to_page_slots.data_ptr(), self.buffer_req_to_page_slots.stride(0),
            req_indices.data_ptr(), context_lens.data_ptr(), s.scores.data_ptr(), s.scores.stride(0),
            batch, int(num_heads), int(num_kv_heads), n_cols, prev_budget, s.packed.data_ptr(), s.packed.stride(0),
            s.lens.data_ptr(), None, s.tickets.data_ptr(), _stream(q)), "svb200_quest_score_select_pack")
        return s.packed[:, : (prev_budget + 1) * page], s.arange, s.lens  # same triple as quest.py:1912-1913
