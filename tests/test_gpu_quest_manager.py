"""Quest decode path through the provider's cache-manager hooks (prepare_decode_static -> per layer
save_rope_kv_if_needed -> build_decode_view -> run_decode) against the oracle restatement of the
reference, step by step across page boundaries.  The manager that hands the
step's K/V rows to the scoring launch and the one that stores them with a launch of their own must
leave bit-identical state."""
from __future__ import annotations

from types import SimpleNamespace

import pytest
import torch

from oracle import attention as oattn
from oracle import quest as oquest

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
PAGE = 16
SKIP = 1  # layer 0 stays dense (quest.py:1816-1821): its store is never deferred


def _build(fuse, *, layers, hkv, bsz, ctx0, max_len, seed):
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.cache_manager.quest import QuestCacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context
    reset_context()
    cfg = configs.make_sparse_config(sparse_method="quest", sink_keep_tokens=16, decode_keep_tokens=96,
                                     recent_keep_tokens=16, quest_skip_layers=SKIP)
    pages_per_row = (max_len + PAGE - 1) // PAGE
    num_pages = bsz * pages_per_row + 3
    mgr = QuestCacheManager(cfg, num_layers=layers, num_kv_heads=hkv, num_pages=num_pages, max_rows=bsz,
                            max_model_len=max_len, device=torch.device(DEV), fuse_decode_store=fuse)
    g = torch.Generator().manual_seed(seed)
    mgr.free_pages_cpu_stack[:] = torch.randperm(num_pages, generator=g).numpy().astype("int32")
    set_context(True)
    for b in range(bsz):
        n = ctx0 - 5 * b
        slots = mgr.allocate_prefill(b, n)
        for layer in range(layers):
            k = torch.randn(n, hkv, 128, generator=g).to(torch.bfloat16).to(DEV)
            v = torch.randn(n, hkv, 128, generator=g).to(torch.bfloat16).to(DEV)
            mgr.save_rope_kv_if_needed(layer, k, v, slots)
    set_context(False, is_long_text=True)
    mgr.set_decode_static_max_context_len(max_len)
    return mgr, cfg


@pytest.mark.parametrize("one_launch", [False, True])
@pytest.mark.parametrize("hq,hkv", [(8, 2), (32, 8)])
def test_quest_decode_steps_through_manager(hq, hkv, one_launch):
    """Two managers step side by side: [0] stores the step's K/V inside the scoring launch -- and, with one_launch,
    runs the whole sparse layer (store, score, select, pack, attention) as ONE launch; [1] uses the stand-alone
    store kernel and two launches."""
    from sparse_vllm_b200.attention_backend import B200AttentionBackend, decode_attention_layer
    from sparse_vllm_b200.cache_manager.base import SparseSelection
    layers, bsz, ctx0, max_len, steps = 3, 3, 1000, 1280, 22
    mgrs = [_build(f, layers=layers, hkv=hkv, bsz=bsz, ctx0=ctx0, max_len=max_len, seed=3)[0] for f in (True, False)]
    cfg = mgrs[0].config
    mgrs[0].fuse_decode_layer = one_launch
    mgrs[1].fuse_decode_layer = False
    backends = [B200AttentionBackend(provider=m._get_provider(hq)) for m in mgrs]
    seqs = [SimpleNamespace(seq_id=b, decode_input_token=1, decode_input_position=0) for b in range(bsz)]
    bufs = [SimpleNamespace(ids=torch.zeros(bsz, dtype=torch.int64, device=DEV),
                            pos=torch.zeros(bsz, dtype=torch.int64, device=DEV),
                            slot=torch.zeros(bsz, dtype=torch.int32, device=DEV),
                            ctx=torch.zeros(bsz, dtype=torch.int32, device=DEV),
                            req=torch.zeros(bsz, dtype=torch.int32, device=DEV)) for _ in mgrs]
    gen = torch.Generator().manual_seed(11)
    fused_launches = 0
    for step in range(steps):
        q = torch.randn(layers, bsz, hq, 128, generator=gen).to(torch.bfloat16)
        k_new = (torch.randn(layers, bsz, hkv, 128, generator=gen) * 2).to(torch.bfloat16)
        v_new = torch.randn(layers, bsz, hkv, 128, generator=gen).to(torch.bfloat16)
        for m, bf in zip(mgrs, bufs):
            m.prepare_decode_static(seqs, bf.ids, bf.pos, bf.slot, bf.ctx, bf.req)
        assert torch.equal(bufs[0].slot, bufs[1].slot) and torch.equal(bufs[0].ctx, bufs[1].ctx)
        for layer in range(layers):
            outs = []
            for m, be, bf in zip(mgrs, backends, bufs):
                m.save_rope_kv_if_needed(layer, k_new[layer].to(DEV), v_new[layer].to(DEV))
                st = m.layer_batch_state
                sel = SparseSelection(kind="full", req_indices=st.req_indices, context_lens=st.context_lens,
                                      max_context_len=st.max_context_len)
                outs.append(decode_attention_layer(m, be, layer, q[layer].to(DEV), sel, num_heads=hq, num_kv_heads=hkv))
            torch.cuda.synchronize()
            fused, plain = mgrs
            # ---- fused store == stand-alone store, bit for bit (one launch: another CTA cut of the attention, so
            # its fp32 partial sums meet in another order)
            if one_launch:
                torch.testing.assert_close(outs[0].float(), outs[1].float(), atol=2e-2, rtol=1e-2)
            else:
                assert torch.equal(outs[0], outs[1])
            assert torch.equal(fused.kv_cache[:, layer], plain.kv_cache[:, layer])
            assert torch.equal(fused.bounds[layer], plain.bounds[layer])
            if layer < SKIP:  # dense: the whole row through the token table
                k_cache, v_cache = (t.cpu() for t in fused.get_layer_kv_cache(layer))
                ctx = bufs[0].ctx.cpu()
                ref = oattn.gqa_decode(q[layer], k_cache, v_cache, fused.buffer_req_to_token_slots.cpu(), bufs[0].req.cpu(),
                                       ctx, max_len_in_batch=int(ctx.max()))
                torch.testing.assert_close(outs[0].cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
                continue
            sf, sp = fused._scratch[bsz], plain._scratch[bsz]
            assert torch.equal(sf.packed, sp.packed) and torch.equal(sf.lens, sp.lens)
            # ---- against the oracle on the state the manager holds
            k_cache, v_cache = (t.cpu() for t in fused.get_layer_kv_cache(layer))
            slots, ctx = bufs[0].slot.cpu().long(), bufs[0].ctx.cpu()
            assert torch.equal(k_cache[slots], k_new[layer]) and torch.equal(v_cache[slots], v_new[layer])
            mx, mn = oquest.build_bounds_from_cache(k_cache)
            page_table = fused.buffer_req_to_page_slots.cpu()
            by_page = fused.bounds_by_page(layer).cpu()
            for b in range(bsz):  # bounds of every completed page, incl. the one this step may have completed
                done = page_table[b, : int(ctx[b]) // PAGE].long()
                assert torch.equal(by_page[done, 0], mx[done])
                assert torch.equal(by_page[done, 1], mn[done])
            view = oquest.build_decode_view_static(
                q[layer], mx, mn, page_table, fused.buffer_req_to_token_slots.cpu(), bufs[0].req.cpu(), ctx,
                token_budget=int(cfg.quest_token_budget), num_kv_heads=hkv, max_context_len=max_len,
                max_pages_per_row=fused.max_pages_per_row, is_long_text=True)
            width = view["packed_slots"].shape[1]
            assert torch.equal(sf.lens.cpu(), view["lens"])
            prev = width // PAGE - 1
            got_pages = sf.packed.cpu()[:, :width:PAGE] // PAGE  # physical page of every selected page
            for b in range(bsz):
                inv = {int(p): i for i, p in enumerate(page_table[b].tolist()) if p >= 0}
                got = [inv[int(p)] for p in got_pages[b, :prev]]
                n_valid = (int(ctx[b]) + PAGE - 1) // PAGE - 1
                # a one-ulp score difference may swap pages, but only pages within one ulp of the k-th score
                c = oquest.classify_selection_ulp(view["scores"][b], view["selected_pages"][b, :prev].tolist(), got,
                                                  n_valid, ulp_slack=1)
                assert c["bad"] == 0, (step, layer, b, c["detail"])
                d = oquest.bf16_ulp_distance(sf.scores[b, :n_valid].cpu(), view["scores"][b, :n_valid])
                assert int(d.max()) <= 1, (step, layer, b)
                assert inv[int(got_pages[b, prev])] == n_valid  # the last page rides along
            ref = oattn.gqa_decode(q[layer], k_cache, v_cache, sf.packed.cpu()[:, :width], torch.arange(bsz, dtype=torch.int32),
                                   sf.lens.cpu(), max_len_in_batch=width)
            torch.testing.assert_close(outs[0].cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
        for m in mgrs:
            m.on_forward_end(seqs, is_prefill=False)
    stats = backends[0].provider.runtime_kernel_stats()["kernel_paths"]
    path = "quest_decode_layer_one_launch" if one_launch else "quest_store_score_select_pack_fused"
    fused_launches = stats.get(path, {}).get("launches", 0)
    assert fused_launches == steps * (layers - SKIP), stats
    assert "quest_store_score_select_pack_fused" not in backends[1].provider.runtime_kernel_stats()["kernel_paths"]
