"""Host-side logic of the Quest cache manager on CPU tensors, launches replaced by recorders: slot / page
allocation of a decode step (quest.py:1280-1360, 1543-1605) and the hand-over of the step's K/V rows to the
layer's scoring launch (when it is deferred, when it is not, who flushes it)."""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np
import pytest
import torch

PAGE = 16


class _Recorder:
    def __init__(self):
        self.calls = []

    def __call__(self, name):
        def fn(*a, **k):
            self.calls.append((name, a, k))
        return fn

    def names(self):
        return [c[0] for c in self.calls]


class _FakeProvider:
    def __init__(self):
        self.select_calls = []

    def quest_select(self, q, bounds_layer, unit_table, page_table, req_rows, context_lens, **kw):
        self.select_calls.append(kw)


@pytest.fixture()
def rig(monkeypatch):
    from sparse_vllm_b200 import configs, ops
    from sparse_vllm_b200.cache_manager.base import CacheManager
    from sparse_vllm_b200.cache_manager.quest import QuestCacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context
    rec = _Recorder()
    monkeypatch.setattr(CacheManager, "_allow_host_tensors", True)
    monkeypatch.setattr(ops, "store_kv_minmax", rec("store_kv_minmax"))
    monkeypatch.setattr(ops, "quest_page_minmax", rec("quest_page_minmax"))
    monkeypatch.setattr(ops, "sched_scratch", lambda rows, device: torch.zeros(1 + 2 * rows, dtype=torch.int32))
    reset_context()

    def make(**kw):
        cfg = configs.make_sparse_config(sparse_method="quest", sink_keep_tokens=16, decode_keep_tokens=96,
                                         recent_keep_tokens=16, quest_skip_layers=1)
        prov = _FakeProvider()
        mgr = QuestCacheManager(cfg, num_layers=2, num_kv_heads=2, num_pages=64, max_rows=3, max_model_len=320,
                                device=torch.device("cpu"), provider=prov, **kw)
        set_context(True)
        for b, n in enumerate((200, 192, 175)):  # row 1 ends exactly on a page boundary
            slots = mgr.allocate_prefill(b, n)
            for layer in range(2):
                mgr.save_rope_kv_if_needed(layer, torch.zeros(n, 2, 128, dtype=torch.bfloat16),
                                           torch.zeros(n, 2, 128, dtype=torch.bfloat16), slots)
        set_context(False, is_long_text=True)
        mgr.set_decode_static_max_context_len(320)
        rec.calls.clear()
        return mgr, prov

    yield SimpleNamespace(make=make, rec=rec)
    reset_context()


def test_managers_refuse_host_tensors_by_default():
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.cache_manager.quest import QuestCacheManager
    cfg = configs.make_sparse_config(sparse_method="quest", sink_keep_tokens=16, decode_keep_tokens=96, recent_keep_tokens=16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        QuestCacheManager(cfg, num_layers=1, num_kv_heads=2, num_pages=8, max_rows=1, max_model_len=64,
                          device=torch.device("cpu"))


def _step_buffers(n):
    return (torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.int32),
            torch.zeros(n, dtype=torch.int32), torch.zeros(n, dtype=torch.int32))


def test_prepare_decode_static_allocates_slots_and_pads(rig):
    mgr, _ = rig.make()
    free_before = mgr._num_free_pages
    seqs = [SimpleNamespace(seq_id=b, decode_input_token=7 + b, decode_input_position=b) for b in range(3)]
    ids, pos, slot, ctx, req = _step_buffers(4)  # graph batch 4 > real batch 3: one padding entry
    mgr.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    assert ctx.tolist() == [201, 193, 176, 201] and req.tolist() == [0, 1, 2, 0]
    assert slot[3] == -1 and ids.tolist() == [7, 8, 9, 7]
    # row 1 was full to a page boundary: exactly one new page was popped, its slot is the page's first
    assert mgr._num_free_pages == free_before - 1
    pt = mgr.buffer_req_to_page_slots_cpu
    assert int(slot[1]) == int(pt[1, 192 // PAGE]) * PAGE
    assert int(slot[0]) == int(pt[0, 200 // PAGE]) * PAGE + 200 % PAGE
    assert int(mgr.buffer_req_to_token_slots[2, 175]) == int(slot[2])
    assert np.array_equal(mgr.buffer_req_to_page_slots.numpy(), pt)
    assert mgr.layer_batch_state.max_context_len == 320  # the pinned bucket, not max(ctx)


def test_decode_store_is_handed_to_the_scoring_launch(rig):
    from sparse_vllm_b200.cache_manager.base import SparseSelection
    mgr, prov = rig.make()
    seqs = [SimpleNamespace(seq_id=b, decode_input_token=1, decode_input_position=0) for b in range(3)]
    ids, pos, slot, ctx, req = _step_buffers(3)
    mgr.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    k = torch.zeros(3, 2, 128, dtype=torch.bfloat16)
    q = torch.zeros(3, 8, 128, dtype=torch.bfloat16)
    sel = SparseSelection(kind="full", req_indices=req, context_lens=ctx, max_context_len=320)
    # dense layer (< quest_skip_layers): stored at once, with the fused bounds refresh
    mgr.save_rope_kv_if_needed(0, k, k)
    assert rig.rec.names() == ["store_kv_minmax"] and rig.rec.calls[0][1][5] is not None
    mgr.build_decode_compute_view(0, q, sel, num_heads=8, num_kv_heads=2)
    assert prov.select_calls == []
    # sparse layer: nothing is launched by the store hook; the scoring launch receives the rows
    rig.rec.calls.clear()
    mgr.save_rope_kv_if_needed(1, k, k)
    assert rig.rec.calls == [] and mgr._pending_store is not None
    view = mgr.build_decode_compute_view(1, q, sel, num_heads=8, num_kv_heads=2)
    st = prov.select_calls[-1]["store"]
    assert st is not None and st.key is k and st.slot_mapping is slot and st.refresh_bounds
    assert mgr._pending_store is None and rig.rec.calls == []
    assert view.meta.active_slots.shape == (3, (7 + 1) * PAGE) and prov.select_calls[-1]["prev_budget"] == 7
    # whoever looks at the cache before the scoring launch gets the rows stored first
    mgr.save_rope_kv_if_needed(1, k, k)
    mgr.get_layer_kv_cache(1)
    assert rig.rec.names() == ["store_kv_minmax"] and mgr._pending_store is None
    rig.rec.calls.clear()
    mgr.save_rope_kv_if_needed(1, k, k)
    mgr.on_forward_end(seqs, is_prefill=False)
    assert rig.rec.names() == ["store_kv_minmax"]


def test_store_is_not_deferred_when_it_cannot_be_fused(rig):
    from sparse_vllm_b200.utils_context import set_context
    mgr, _ = rig.make(fuse_decode_store=False)
    seqs = [SimpleNamespace(seq_id=b, decode_input_token=1, decode_input_position=0) for b in range(3)]
    ids, pos, slot, ctx, req = _step_buffers(3)
    mgr.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    k = torch.zeros(3, 2, 128, dtype=torch.bfloat16)
    mgr.save_rope_kv_if_needed(1, k, k)
    assert rig.rec.names() == ["store_kv_minmax"]
    mgr2, prov2 = rig.make()
    mgr2.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    set_context(False, is_long_text=False)  # mixed batch: two launches, the store goes first
    mgr2.save_rope_kv_if_needed(1, k, k)
    assert rig.rec.names() == ["store_kv_minmax"] and mgr2._pending_store is None


def test_bounds_chunks_follow_page_allocation(rig):
    """Every (row, logical page // 16) owns one bounds chunk; page_rec maps a physical page to its record;
    a decode token that opens logical page 16 opens a new chunk; freeing a sequence returns its chunks."""
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.cache_manager.quest import QuestCacheManager
    cfg = configs.make_sparse_config(sparse_method="quest", sink_keep_tokens=16, decode_keep_tokens=96, recent_keep_tokens=16)
    mgr = QuestCacheManager(cfg, num_layers=1, num_kv_heads=2, num_pages=64, max_rows=2, max_model_len=512,
                            device=torch.device("cpu"), provider=_FakeProvider())
    assert mgr.num_chunks == 64 // PAGE + 2 and mgr._num_free_chunks == mgr.num_chunks
    assert tuple(mgr.bounds.shape) == (1, mgr.num_chunks, 1, PAGE, 2, 2, 128)
    mgr.allocate_prefill(0, 256)  # exactly 16 pages: one chunk
    mgr.allocate_prefill(1, 300)  # 19 pages: two chunks
    ut, rec, pt = mgr.unit_table_cpu, mgr.page_rec.numpy(), mgr.buffer_req_to_page_slots_cpu
    assert (ut[0] >= 0).sum() == 1 and (ut[1] >= 0).sum() == 2 and mgr._num_free_chunks == mgr.num_chunks - 3
    assert np.array_equal(mgr.unit_table.numpy(), ut)
    for row, n_pages in ((0, 16), (1, 19)):
        j = np.arange(n_pages)
        assert np.array_equal(rec[pt[row, :n_pages]], ut[row, j // PAGE] * PAGE + j % PAGE)
    assert len(set(ut[ut >= 0].tolist())) == 3 and (rec >= 0).sum() == 35
    # the next decode token of row 0 opens logical page 16 -> a second chunk for row 0
    seqs = [SimpleNamespace(seq_id=0, decode_input_token=1, decode_input_position=0)]
    mgr.prepare_decode_static(seqs, *_step_buffers(1))
    assert (ut[0] >= 0).sum() == 2 and int(rec[pt[0, 16]]) == int(ut[0, 1]) * PAGE
    assert np.array_equal(mgr.unit_table.numpy(), ut)
    mgr.free_seq(1)
    assert mgr._num_free_chunks == mgr.num_chunks - 2 and (mgr.unit_table_cpu[1] == -1).all()
    assert (mgr.page_rec.numpy() >= 0).sum() == 17
