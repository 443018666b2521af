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
