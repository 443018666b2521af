"""Host-side logic of the H2O cache manager on CPU tensors, launches replaced by recorders: which rows a burst
evicts (own trigger, pressure branch with unscheduled rows, h2o.py:1498-1540), the offsets / lengths / free-stack
positions handed to the compaction launch, the packed metadata of a decode step's preparation and of the prefill
window scoring (h2o.py:257-476, 665-894)."""
from __future__ import annotations

from types import SimpleNamespace

import numpy as np
import pytest
import torch


class _Rec:
    def __init__(self):
        self.calls = []

    def __call__(self, name):
        def fn(*a, **k):
            self.calls.append((name, a, k))
        return fn


@pytest.fixture()
def rig(monkeypatch):
    from sparse_vllm_b200 import configs, ops
    from sparse_vllm_b200.cache_manager.base import CacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context
    rec = _Rec()
    monkeypatch.setattr(CacheManager, "_allow_host_tensors", True)
    for name in ("store_kv_minmax", "h2o_select_compact", "h2o_normalize_accumulate", "h2o_prefill_score"):
        monkeypatch.setattr(ops, name, rec(name))

    def prepare(meta, free_stack, slot_tables, layer_slot_mapping, context_lens, req_indices, slot_mapping, kv_lens,
                rows_out, batch):
        """What svb200_h2o_prepare_decode does, on the host (the kernel is checked by the GPU manager tests)."""
        m = meta.numpy()
        rows, lens, top = m[:batch], m[batch: 2 * batch], int(m[2 * batch])
        for layer in range(slot_tables.shape[0]):
            for b in range(batch):
                slot = int(free_stack[layer, top - batch + b])
                slot_tables[layer, int(rows[b]), int(lens[b]) - 1] = slot
                layer_slot_mapping[layer, b] = slot
        context_lens[:batch] = torch.from_numpy(lens.copy())
        req_indices[:batch] = torch.from_numpy(rows.copy())
        slot_mapping[:batch] = layer_slot_mapping[0, :batch]
        kv_lens[:batch] = torch.from_numpy(lens.copy())
        rows_out[:batch] = torch.from_numpy(rows.copy())
    monkeypatch.setattr(ops, "h2o_prepare_decode", prepare)
    reset_context()
    set_context(False)

    def make(*, budget=48, interval=16, prefill_budget=48, rows=4, slots=600, layers=2, chunk=0, window=32):
        cfg = configs.make_sparse_config(sparse_method="h2o", h2o_decode_budget=budget, h2o_decode_eviction_interval=interval,
                                         h2o_prefill_budget=prefill_budget, h2o_recent_ratio=0.5, h2o_prefill_score_window=window)
        return CacheManager.create(cfg, num_layers=layers, num_kv_heads=2, num_slots=slots, max_rows=rows,
                                   device=torch.device("cpu"), max_prefill_chunk=chunk, provider=object())
    yield SimpleNamespace(make=make, rec=rec)
    reset_context()


def _admit(mgr, sid, n):
    L = mgr.num_layers
    z = torch.zeros(L, n, 2, 128, dtype=torch.bfloat16)
    return mgr.admit_compacted_prompt(sid, z, z, torch.zeros(L, n))


def _bufs(n):
    return (torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.int64), torch.zeros(n, dtype=torch.int32),
            torch.zeros(n, dtype=torch.int32), torch.zeros(n, dtype=torch.int32))


def test_decode_step_preparation_and_per_row_eviction(rig):
    mgr = rig.make()
    for sid, n in ((0, 30), (1, 48), (2, 41), (3, 48)):
        _admit(mgr, sid, n)
    mgr.free_seq(1)
    seqs = [SimpleNamespace(seq_id=s) for s in (0, 2, 3)]
    ids, pos, slot, ctx, req = _bufs(3)
    free0 = mgr.num_free
    top = mgr.free_stack[:, free0 - 3: free0].clone()
    mgr.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    assert ctx.tolist() == [31, 42, 49] and req.tolist() == [0, 2, 3] and mgr.num_free == free0 - 3
    assert torch.equal(mgr.slot_tables[:, [0, 2, 3], [30, 41, 48]], top)       # the popped slots close the rows
    assert torch.equal(mgr.layer_slot_mapping(1), top[1]) and slot.tolist() == top[0].tolist()
    assert mgr.row_seq_lens[[0, 2, 3]].tolist() == [31, 42, 49]
    # layer end: the batch's cumulative rows are addressed by request row, the step buffer is handed back pre-filled
    rig.rec.calls.clear()
    mgr.on_layer_attention_end(1)
    name, a, k = rig.rec.calls[-1]
    assert name == "h2o_normalize_accumulate" and k["cum_rows"].tolist() == [0, 2, 3] and k["refill"] == -1e20
    assert k["width"] == 64 and k["kv_lens"].tolist() == [31, 42, 49]
    # no row reached budget + interval and slots are left: nothing to evict
    assert mgr.evict_after_decode() is False and mgr.last_evicted_rows == []
    # 15 more steps: row 3 (49 -> 64) reaches budget + interval on its own
    for _ in range(15):
        mgr.prepare_decode_static(seqs, ids, pos, slot, ctx, req)
    rig.rec.calls.clear()
    free1 = mgr.num_free
    assert mgr.evict_after_decode() is True and mgr.last_evicted_rows == [3]
    name, a, k = rig.rec.calls[-1]
    assert name == "h2o_select_compact" and k["kv_len"] == 64 and k["budget"] == 48 and k["recent"] == 24
    W, R, S = mgr.width, mgr.max_rows, mgr.num_slots
    assert a[1].tolist() == [(0 * R + 3) * W, (1 * R + 3) * W] and k["kv_lens"].tolist() == [64, 64]
    assert k["free_rows"].tolist() == [0 * S + free1, 1 * S + free1]
    assert mgr.num_free == free1 + 16 and int(mgr.row_seq_lens[3]) == 48


def test_pressure_evicts_every_active_row_over_budget(rig):
    mgr = rig.make(slots=30 + 48 + 41 + 48 + 11)
    for sid, n in ((0, 30), (1, 48), (2, 41), (3, 48)):
        _admit(mgr, sid, n)
    assert mgr.num_free == 11
    all_seqs = [SimpleNamespace(seq_id=s) for s in range(4)]
    mgr.prepare_decode_static(all_seqs, *_bufs(4))   # 7 left; every row is now an active decode row
    assert mgr.evict_after_decode() is False
    mgr.prepare_decode_static(all_seqs, *_bufs(4))   # 3 left
    assert mgr.evict_after_decode() is False
    three = [SimpleNamespace(seq_id=s) for s in (0, 1, 2)]
    mgr.prepare_decode_static(three, *_bufs(3))      # 0 left: seq 3 sat this step out
    rig.rec.calls.clear()
    assert mgr.evict_after_decode() is True
    # trigger budget + 1: rows 1 (51) and 3 (50, unscheduled) are over it, rows 0 (33) and 2 (44) are not;
    # scheduled rows first, then the other active rows in ascending order (h2o.py:1518-1522)
    assert mgr.last_evicted_rows == [1, 3]
    name, a, k = rig.rec.calls[-1]
    S = mgr.num_slots
    assert k["kv_len"] == 51 and k["kv_lens"].tolist() == [51, 50, 51, 50]
    assert k["free_rows"].tolist() == [0, 3, S, S + 3]          # row 1 drops 3 slots, row 3 the next 2, per layer
    assert mgr.num_free == 5 and mgr.row_seq_lens[[1, 3]].tolist() == [48, 48]
    with pytest.raises(RuntimeError, match="Out of H2O KV slots"):
        for _ in range(3):
            mgr.prepare_decode_static(all_seqs, *_bufs(4))


def test_prefill_chunks_score_ranges_and_budgets(rig):
    mgr = rig.make(budget=64, interval=64, prefill_budget=128, chunk=96, slots=2 * 224 + 8, rows=2, window=32)
    assert mgr.width == 224 and mgr.decode_width == 128 and tuple(mgr.step_scores.shape) == (2, 2, 128)
    s0 = mgr.allocate_prefill_chunk(5, 96)
    s1 = mgr.allocate_prefill_chunk(6, 20)
    assert tuple(s0.shape) == (2, 96) and torch.equal(mgr.slot_tables[:, 0, :96], s0) and torch.equal(mgr.slot_tables[:, 1, :20], s1)
    q = torch.zeros(116, 8, 128, dtype=torch.bfloat16)
    rig.rec.calls.clear()
    mgr.collect_prefill_attention_score(1, q, [5, 6], [96, 20])
    name, a, k = rig.rec.calls[-1]
    assert name == "h2o_prefill_score"
    req, start_loc, ctx, cache, starts, ends = (t.tolist() for t in a[3:9])
    assert (req, start_loc, ctx, cache, starts, ends) == ([0, 1], [0, 96], [96, 20], [0, 0], [64, 0], [96, 20])
    assert k["max_query_len"] == 32 and k["max_context_len"] == 96 and k["cum_rows"].tolist() == [0, 1]
    assert mgr.evict_after_prefill([5, 6], final=False) is False          # nobody over the prefill budget yet
    mgr.allocate_prefill_chunk(5, 96)                                       # 192 > 128
    rig.rec.calls.clear()
    mgr.collect_prefill_attention_score(0, q[:96], [5], [96])
    a = rig.rec.calls[-1][1]
    assert [t.tolist() for t in a[5:9]] == [[192], [96], [160], [192]]     # ctx, cached prefix, window start / end
    assert mgr.evict_after_prefill([5, 6], final=False) is True and mgr.last_evicted_rows == [0]
    k = rig.rec.calls[-1][2]
    assert k["budget"] == 128 and k["recent"] == 64 and k["kv_lens"].tolist() == [192, 192]
    assert int(mgr.row_seq_lens[0]) == 128
    assert mgr.evict_after_prefill([5, 6], final=True) is True and mgr.last_evicted_rows == [0]   # final: decode budget
    assert rig.rec.calls[-1][2]["budget"] == 64 and mgr.row_seq_lens.tolist() == [64, 20]
    with pytest.raises(RuntimeError, match="does not fit the row"):
        mgr.allocate_prefill_chunk(5, 200)
