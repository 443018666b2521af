"""H2O decode path through the provider's cache-manager hooks (scored decode, layer-end normalise +
accumulate, burst eviction) against the oracle restatement of the reference, step by step over one
full eviction interval."""
from __future__ import annotations

import math
from types import SimpleNamespace

import pytest
import torch

from oracle import attention as oattn
from oracle import h2o as oh2o

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _make(budget, interval, layers, hkv, bsz, slots):
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.cache_manager.base import CacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context
    reset_context()
    set_context(False)
    cfg = configs.make_sparse_config(sparse_method="h2o", h2o_decode_budget=budget,
                                     h2o_decode_eviction_interval=interval, h2o_prefill_budget=budget,
                                     h2o_recent_ratio=0.5)
    mgr = CacheManager.create(cfg, num_layers=layers, num_kv_heads=hkv, num_slots=slots, max_rows=bsz, device=DEV)
    assert type(mgr).__name__ == "H2OCacheManager"
    return mgr


@pytest.mark.parametrize("hq,hkv", [(8, 2), (28, 4)])
def test_h2o_decode_interval_and_eviction(hq, hkv):
    from sparse_vllm_b200.attention_backend import B200AttentionBackend, decode_attention_layer
    budget, interval, layers, bsz = 64, 64, 2, 3
    width = budget + interval
    mgr = _make(budget, interval, layers, hkv, bsz, slots=bsz * width + 32)
    mgr.refill_in_kernel = False  # this test looks at the normalised step scores (the reference's in-place softmax)
    gen = torch.Generator().manual_seed(5 + hq)
    for b in range(bsz):
        k = torch.randn(layers, budget, hkv, 128, generator=gen).to(torch.bfloat16)
        v = torch.randn(layers, budget, hkv, 128, generator=gen).to(torch.bfloat16)
        s = torch.rand(layers, budget, generator=gen)
        assert mgr.admit_compacted_prompt(b, k.to(DEV), v.to(DEV), s.to(DEV)) == b
    backend = B200AttentionBackend(provider=mgr._get_provider(hq))
    seqs = [SimpleNamespace(seq_id=b) for b in range(bsz)]
    ids = torch.zeros(bsz, dtype=torch.int64, device=DEV)
    pos = torch.zeros(bsz, dtype=torch.int64, device=DEV)
    slot_mapping = torch.zeros(bsz, dtype=torch.int32, device=DEV)
    ctx = torch.zeros(bsz, dtype=torch.int32, device=DEV)
    req = torch.zeros(bsz, dtype=torch.int32, device=DEV)
    # oracle-side cumulative scores, rebuilt from the manager's state after every eviction
    cum_ref = mgr.cumulative.view(layers, bsz, width).cpu().clone()
    evictions = 0
    for step in range(interval + 5):
        kv_len = int(mgr.row_seq_lens[0]) + 1
        q = torch.randn(layers, bsz, hq, 128, generator=gen).to(torch.bfloat16)
        k_new = torch.randn(layers, bsz, hkv, 128, generator=gen).to(torch.bfloat16)
        v_new = torch.randn(layers, bsz, hkv, 128, generator=gen).to(torch.bfloat16)
        mgr.prepare_decode_static(seqs, ids, pos, slot_mapping, ctx, req)
        assert ctx.tolist() == [kv_len] * bsz
        for layer in range(layers):
            mgr.save_rope_kv_if_needed(layer, k_new[layer].to(DEV), v_new[layer].to(DEV))
            sel = mgr.decode_selection(layer)
            out = decode_attention_layer(mgr, backend, layer, q[layer].to(DEV), sel, num_heads=hq, num_kv_heads=hkv)
            torch.cuda.synchronize()
            # ---- oracle on the same state
            k_cache, v_cache = (t.cpu() for t in mgr.get_layer_kv_cache(layer))
            table = mgr.slot_tables[layer].cpu()
            raw = torch.full((bsz, width), oh2o.SCORE_FILL, dtype=torch.float32)
            ref = oattn.gqa_decode(q[layer], k_cache, v_cache, table, req.cpu(), ctx.cpu(), max_len_in_batch=width,
                                   attn_score=raw)
            torch.testing.assert_close(out.cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
            norm = oh2o.normalize_scores_(raw.clone(), 1.0 / math.sqrt(128))
            got_step = mgr.step_scores[layer, :bsz].cpu()
            torch.testing.assert_close(got_step, norm, atol=2e-5, rtol=2e-3)
            new_cum = oh2o.accumulate_all_layers(norm[None], cum_ref[layer][None, :, : kv_len - 1], kv_len)[0]
            cum_ref[layer, :, :kv_len] = new_cum
            torch.testing.assert_close(mgr.cumulative.view(layers, bsz, width)[layer, :, :kv_len].cpu(), new_cum,
                                       atol=5e-5, rtol=2e-3)
        # ---- eviction: selection is bit-exact w.r.t. the scores the GPU holds
        before_cum = mgr.cumulative.view(layers, bsz, width).cpu().clone()
        before_tab = mgr.slot_tables.cpu().clone()
        free_before = mgr.num_free
        keep = torch.zeros(layers * bsz, budget, dtype=torch.int32, device=DEV)
        if mgr.evict_after_decode(keep_out=keep):
            evictions += 1
            assert kv_len == budget + interval
            torch.cuda.synchronize()
            sc = before_cum[:, :, :kv_len].reshape(layers * bsz, kv_len)
            ref_keep = oh2o.select_indices_batch(sc, budget=budget, recent_ratio=0.5)
            assert torch.equal(keep.cpu().long(), ref_keep)
            exp_tab, exp_sc, exp_drop = oh2o.compact_rows(before_tab[:, :, :kv_len].reshape(layers * bsz, kv_len), sc, ref_keep)
            assert torch.equal(mgr.slot_tables[:, :, :kv_len].cpu().reshape(layers * bsz, kv_len), exp_tab)
            assert torch.equal(mgr.cumulative.view(layers, bsz, width)[:, :, :budget].cpu().reshape(layers * bsz, budget), exp_sc)
            drop = kv_len - budget
            assert mgr.num_free == free_before + bsz * drop and int(mgr.row_seq_lens[0]) == budget
            pushed = mgr.free_stack[:, free_before: free_before + bsz * drop].cpu().reshape(layers * bsz, drop)
            assert torch.equal(pushed, exp_drop)
            cum_ref = mgr.cumulative.view(layers, bsz, width).cpu().clone()
    assert evictions == 1


@pytest.mark.parametrize("tight", [False, True])
def test_h2o_ragged_rows_evict_on_their_own_trigger(tight):
    """Rows of different length on non-consecutive request rows (h2o.py:1498-1625): every row evicts when IT
    reaches budget + interval; with the free stack empty (``tight``) every active row over budget evicts at
    once (the pressure branch, 1510-1522), including a row that was not scheduled in the step.  Cumulative
    scores follow the oracle step by step; the step-score rows come back pre-filled for the next step."""
    from sparse_vllm_b200.attention_backend import B200AttentionBackend, decode_attention_layer
    hq, hkv, budget, interval, layers = 8, 2, 48, 16, 2
    width = budget + interval
    prompt = {0: 30, 1: 48, 2: 41, 3: 48}
    # tight: exactly enough slots that the stack runs dry before any row reaches budget + interval
    slots = (30 + 41 + 48) + 3 * 9 if tight else 4 * width + 8
    mgr = _make(budget, interval, layers, hkv, 4, slots=slots + 64)
    gen = torch.Generator().manual_seed(17)
    for sid, n in prompt.items():
        k = torch.randn(layers, n, hkv, 128, generator=gen).to(torch.bfloat16)
        v = torch.randn(layers, n, hkv, 128, generator=gen).to(torch.bfloat16)
        mgr.admit_compacted_prompt(sid, k.to(DEV), v.to(DEV), torch.rand(layers, n, generator=gen).to(DEV))
    mgr.free_seq(1)  # rows 0, 2, 3 stay: not consecutive
    if tight:
        mgr.num_free = 3 * 8 + 2  # the allocator's view of the stack: it runs dry in step 8 (two rows scheduled)
    backend = B200AttentionBackend(provider=mgr._get_provider(hq))
    sids = [0, 2, 3]
    rows = [mgr.seq_id_to_row[s] for s in sids]
    assert rows != list(range(rows[0], rows[0] + 3))
    bufs = dict(ids=torch.zeros(3, dtype=torch.int64, device=DEV), pos=torch.zeros(3, dtype=torch.int64, device=DEV),
                slot=torch.zeros(3, dtype=torch.int32, device=DEV), ctx=torch.zeros(3, dtype=torch.int32, device=DEV),
                req=torch.zeros(3, dtype=torch.int32, device=DEV))
    cum_ref = mgr.cumulative.view(layers, 4, width).cpu().clone()
    evicted_at = {}
    for step in range(12 if tight else 45):  # tight: the slots the burst frees last three more steps
        sched = sids if not (tight and step == 8) else sids[:2]  # tight: seq 3 sits out the step that runs dry
        n = len(sched)
        seqs = [SimpleNamespace(seq_id=s) for s in sched]
        rws = [mgr.seq_id_to_row[s] for s in sched]
        lens = [int(mgr.row_seq_lens[r]) + 1 for r in rws]
        q = torch.randn(layers, n, hq, 128, generator=gen).to(torch.bfloat16)
        kn = torch.randn(layers, n, hkv, 128, generator=gen).to(torch.bfloat16)
        vn = torch.randn(layers, n, hkv, 128, generator=gen).to(torch.bfloat16)
        mgr.prepare_decode_static(seqs, bufs["ids"][:n], bufs["pos"][:n], bufs["slot"][:n], bufs["ctx"][:n], bufs["req"][:n])
        assert bufs["ctx"][:n].tolist() == lens and bufs["req"][:n].tolist() == rws
        for layer in range(layers):
            mgr.save_rope_kv_if_needed(layer, kn[layer].to(DEV), vn[layer].to(DEV))
            out = decode_attention_layer(mgr, backend, layer, q[layer].to(DEV), mgr.decode_selection(layer), num_heads=hq,
                                         num_kv_heads=hkv)
            torch.cuda.synchronize()
            k_cache, v_cache = (t.cpu() for t in mgr.get_layer_kv_cache(layer))
            raw = torch.full((n, width), oh2o.SCORE_FILL, dtype=torch.float32)
            ref = oattn.gqa_decode(q[layer], k_cache, v_cache, mgr.slot_tables[layer].cpu(), bufs["req"][:n].cpu(),
                                   bufs["ctx"][:n].cpu(), max_len_in_batch=width, attn_score=raw)
            torch.testing.assert_close(out.cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
            norm = oh2o.normalize_scores_(raw.clone(), 1.0 / math.sqrt(128))
            assert bool((mgr.step_scores[layer, :n] == oh2o.SCORE_FILL).all()), "step rows must come back pre-filled"
            for b, (r, kv_len) in enumerate(zip(rws, lens)):
                new = oh2o.accumulate_all_layers(norm[None, b: b + 1], cum_ref[layer, r][None, None, : kv_len - 1], kv_len)[0, 0]
                cum_ref[layer, r, :kv_len] = new
                torch.testing.assert_close(mgr.cumulative.view(layers, 4, width)[layer, r, :kv_len].cpu(), new,
                                           atol=5e-5, rtol=2e-3)
        before_cum = mgr.cumulative.view(layers, 4, width).cpu().clone()
        before_tab = mgr.slot_tables.cpu().clone()
        before_lens = mgr.row_seq_lens.copy()
        free_before = mgr.num_free
        keep = torch.full((layers * 4, budget), -1, dtype=torch.int32, device=DEV)
        if mgr.evict_after_decode(keep_out=keep):
            torch.cuda.synchronize()
            ev = mgr.last_evicted_rows
            dropped = 0
            for j, r in enumerate(ev):
                kv_len = int(before_lens[r])
                evicted_at.setdefault(r, []).append((step, kv_len))
                assert kv_len >= (budget + 1 if free_before <= 0 else budget + interval)
                for layer in range(layers):
                    sc = before_cum[layer, r, :kv_len][None]
                    ref_keep = oh2o.select_indices_batch(sc, budget=budget, recent_ratio=0.5)
                    assert torch.equal(keep[layer * len(ev) + j].cpu().long(), ref_keep[0])
                    exp_tab, exp_sc, exp_drop = oh2o.compact_rows(before_tab[layer, r, :kv_len][None], sc, ref_keep)
                    assert torch.equal(mgr.slot_tables[layer, r, :kv_len].cpu()[None], exp_tab)
                    assert torch.equal(mgr.cumulative.view(layers, 4, width)[layer, r, :budget].cpu()[None], exp_sc)
                    pushed = mgr.free_stack[layer, free_before + dropped: free_before + dropped + kv_len - budget].cpu()
                    assert torch.equal(pushed[None], exp_drop)
                dropped += kv_len - budget
                assert int(mgr.row_seq_lens[r]) == budget
            assert mgr.num_free == free_before + dropped
            cum_ref = mgr.cumulative.view(layers, 4, width).cpu().clone()
    if tight:
        # the step that consumed the last slot evicted EVERY active row over budget, also the unscheduled one
        first = min(v[0][0] for v in evicted_at.values())
        assert first == 8 and sorted(r for r, v in evicted_at.items() if v[0][0] == first) == [rows[1], rows[2]]
        assert all(kv < budget + interval for v in evicted_at.values() for _, kv in v[:1])
    else:
        # each row evicts when IT reaches budget + interval: the longer prompts first
        assert all(kv == budget + interval for v in evicted_at.values() for _, kv in v)
        assert evicted_at[rows[2]][0][0] < evicted_at[rows[1]][0][0] < evicted_at[rows[0]][0][0]


def test_h2o_chunked_prefill_scores_and_compacts_then_decodes():
    """The prefill side through the manager (h2o.py:665-894, 1352-1496): chunks are appended to
    the compressed row of every layer, the last `window` queries of each chunk score the row, cumulative scores
    accumulate with weight = scoring queries, rows over the prefill budget compact after intermediate chunks and
    to the decode budget after the final one; the decode path then runs on what prefill left."""
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.attention_backend import B200AttentionBackend, decode_attention_layer
    from sparse_vllm_b200.cache_manager.base import CacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context
    hq, hkv, layers = 8, 2, 2
    budget, interval, prefill_budget, window, chunk = 64, 64, 128, 32, 96
    reset_context()
    set_context(True)
    cfg = configs.make_sparse_config(sparse_method="h2o", h2o_decode_budget=budget, h2o_decode_eviction_interval=interval,
                                     h2o_prefill_budget=prefill_budget, h2o_recent_ratio=0.5, h2o_prefill_score_window=window)
    mgr = CacheManager.create(cfg, num_layers=layers, num_kv_heads=hkv, num_slots=2 * (prefill_budget + chunk) + 64,
                              max_rows=2, device=DEV, max_prefill_chunk=chunk)
    W = mgr.width
    assert W == prefill_budget + chunk
    gen = torch.Generator().manual_seed(3)
    prompts = {10: [96, 96, 64], 11: [96, 50]}  # chunk lengths per sequence; the last one is the final chunk
    cum_ref = {sid: [None] * layers for sid in prompts}  # oracle-side cumulative vectors per layer
    for step in range(3):
        active = [sid for sid, ch in prompts.items() if step < len(ch)]
        lens = [prompts[sid][step] for sid in active]
        final = {sid: step == len(prompts[sid]) - 1 for sid in active}
        slots = {sid: mgr.allocate_prefill_chunk(sid, n) for sid, n in zip(active, lens)}
        q = torch.randn(layers, sum(lens), hq, 128, generator=gen).to(torch.bfloat16)
        for layer in range(layers):
            off = 0
            for sid, n in zip(active, lens):
                k = torch.randn(n, hkv, 128, generator=gen).to(torch.bfloat16)
                v = torch.randn(n, hkv, 128, generator=gen).to(torch.bfloat16)
                mgr.save_rope_kv_if_needed(layer, k.to(DEV), v.to(DEV), slots[sid][layer].contiguous())
                off += n
            got = mgr.collect_prefill_attention_score(layer, q[layer].to(DEV), active, lens)
            torch.cuda.synchronize()
            # ---- oracle on the manager's state
            k_cache = mgr.get_layer_kv_cache(layer)[0].cpu()
            rows = [mgr.seq_id_to_row[s] for s in active]
            ctx = [int(mgr.row_seq_lens[r]) for r in rows]
            cache = [c - n for c, n in zip(ctx, lens)]
            starts = [max(c, e - window) for c, e in zip(cache, ctx)]
            start_loc = [sum(lens[:i]) for i in range(len(lens))]
            ref = oh2o.prefill_window_score(q[layer], k_cache, mgr.slot_tables[layer].cpu(), torch.tensor(rows, dtype=torch.int32),
                                            torch.tensor(start_loc, dtype=torch.int32), torch.tensor(ctx, dtype=torch.int32),
                                            torch.tensor(cache, dtype=torch.int32), torch.tensor(starts, dtype=torch.int32),
                                            torch.tensor(ctx, dtype=torch.int32))
            for b, (sid, r) in enumerate(zip(active, rows)):
                torch.testing.assert_close(got[b, : ctx[b]].cpu(), ref[b, : ctx[b]], atol=2e-5, rtol=2e-3)
                cum_ref[sid][layer] = oh2o.accumulate_prefill(cum_ref[sid][layer], ref[b], new_len=ctx[b],
                                                              weight=float(ctx[b] - starts[b]))
                torch.testing.assert_close(mgr.cumulative.view(layers, 2, W)[layer, r, : ctx[b]].cpu(), cum_ref[sid][layer],
                                           atol=2e-4, rtol=2e-3)
        # ---- eviction after the chunk: intermediate rows to the prefill budget, final rows to the decode budget
        for is_final in (False, True):
            group = [sid for sid in active if final[sid] == is_final]
            if not group:
                continue
            target = budget if is_final else prefill_budget
            before = {sid: int(mgr.row_seq_lens[mgr.seq_id_to_row[sid]]) for sid in group}
            before_cum = mgr.cumulative.view(layers, 2, W).cpu().clone()
            before_tab = mgr.slot_tables.cpu().clone()
            keep = torch.full((layers * len(group), target), -1, dtype=torch.int32, device=DEV)
            ran = mgr.evict_after_prefill(group, final=is_final, keep_out=keep)
            torch.cuda.synchronize()
            over = [sid for sid in group if before[sid] > target]
            assert ran == bool(over)
            for j, r in enumerate(mgr.last_evicted_rows):
                sid = next(s for s in group if mgr.seq_id_to_row[s] == r)
                n = before[sid]
                for layer in range(layers):
                    sc = before_cum[layer, r, :n][None]
                    ref_keep = oh2o.select_indices_batch(sc, budget=target, recent_ratio=0.5)
                    assert torch.equal(keep[layer * len(mgr.last_evicted_rows) + j].cpu().long(), ref_keep[0])
                    exp_tab, exp_sc, _ = oh2o.compact_rows(before_tab[layer, r, :n][None], sc, ref_keep)
                    assert torch.equal(mgr.slot_tables[layer, r, :n].cpu()[None], exp_tab)
                    assert torch.equal(mgr.cumulative.view(layers, 2, W)[layer, r, :target].cpu()[None], exp_sc)
                    cum_ref[sid][layer] = exp_sc[0]
                assert int(mgr.row_seq_lens[r]) == target
    assert all(int(mgr.row_seq_lens[mgr.seq_id_to_row[s]]) <= budget for s in prompts)
    # ---- the decode path takes over from the state prefill left
    set_context(False)
    backend = B200AttentionBackend(provider=mgr._get_provider(hq))
    seqs = [SimpleNamespace(seq_id=s) for s in prompts]
    bufs = [torch.zeros(2, dtype=dt, device=DEV) for dt in (torch.int64, torch.int64, torch.int32, torch.int32, torch.int32)]
    mgr.prepare_decode_static(seqs, *bufs)
    qd = torch.randn(2, hq, 128, generator=gen).to(torch.bfloat16)
    kd = torch.randn(2, hkv, 128, generator=gen).to(torch.bfloat16)
    mgr.save_rope_kv_if_needed(0, kd.to(DEV), kd.to(DEV))
    out = decode_attention_layer(mgr, backend, 0, qd.to(DEV), mgr.decode_selection(0), num_heads=hq, num_kv_heads=hkv)
    torch.cuda.synchronize()
    k_cache, v_cache = (t.cpu() for t in mgr.get_layer_kv_cache(0))
    ref = oattn.gqa_decode(qd, k_cache, v_cache, mgr.slot_tables[0].cpu(), bufs[4].cpu(), bufs[3].cpu(),
                           max_len_in_batch=mgr.decode_width)
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
