#!/usr/bin/env python
"""Generate tests/golden/*.npz by running the UNMODIFIED reference on CPU.

Run in the build container only (needs /root/reference):

    python oracle/gen_golden.py

Imports ``sparsevllm`` from /root/reference/src with SPARSEVLLM_PLATFORM=cpu, builds the cache
managers with the reference tests' ``object.__new__`` fixture pattern
(/root/reference/tests/test_prefix_cache.py:208-235), and executes the reference's own
functions: QuestCacheManager.on_kv_stored / on_forward_end / build_decode_view,
H2OCacheManager.select_h2o_indices(_batch), SparseController-style normalisation, and the
Triton decode kernels under TRITON_INTERPRET=1 (inputs upcast to fp32: the interpreter is
broken for bf16).  bf16 tensors are stored as their uint16 bit patterns.
"""
from __future__ import annotations

import os
import sys
from pathlib import Path
from types import SimpleNamespace

os.environ.setdefault("SPARSEVLLM_PLATFORM", "cpu")
os.environ.setdefault("CUDA_VISIBLE_DEVICES", "")
os.environ.setdefault("TRITON_INTERPRET", "1")
REF_SRC = "/root/reference/src"
sys.path.insert(0, REF_SRC)

import numpy as np  # noqa: E402
import torch  # noqa: E402

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from oracle import cases  # noqa: E402
GOLDEN = ROOT / "tests" / "golden"
PAGE = 16


def bits(t: torch.Tensor) -> np.ndarray:
    """bf16 tensor -> uint16 bit pattern array."""
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def make_quest_manager(*, layers, hkv, d, num_pages, rows, max_model_len, token_budget, skip_layers=0):
    from sparsevllm.engine.cache_manager.quest import QuestCacheManager

    m = object.__new__(QuestCacheManager)
    m.config = SimpleNamespace(quest_skip_layers=skip_layers, quest_token_budget=token_budget)
    m.runtime_layout = SimpleNamespace(kv_layer_index=lambda layer_idx: int(layer_idx))
    m.device = torch.device("cpu")
    m.enable_prefix_caching = False
    m.page_size = PAGE
    m.num_pages = num_pages
    m.num_kv_heads = hkv
    m.head_dim = d
    m.num_kv_layers = layers
    m.max_model_len = max_model_len
    m.max_pages_per_row = (max_model_len + PAGE - 1) // PAGE
    m.page_offsets_i32 = torch.arange(PAGE, dtype=torch.int32)
    m.page_offsets_i64 = m.page_offsets_i32.to(torch.int64)
    m.kv_cache = torch.zeros(2, layers, num_pages * PAGE, hkv, d, dtype=torch.bfloat16)
    m.metadata_cache = torch.zeros(2, layers, num_pages, hkv, d, dtype=torch.bfloat16)
    m.buffer_req_to_token_slots = torch.zeros((rows, max_model_len), dtype=torch.int32)
    m.buffer_req_to_page_slots = torch.full((rows, m.max_pages_per_row), -1, dtype=torch.int32)
    m.buffer_req_to_page_slots_cpu = m.buffer_req_to_page_slots.numpy()
    m.layer_batch_state = SimpleNamespace(max_context_len=None)
    m.row_seq_lens = np.zeros((rows,), dtype=np.int32)
    m.seq_id_to_row = {}
    m._prefill_metadata_full_pages = False
    m.prefix_offload_controller = None
    return m


def quest_case(name, *, seed, bsz, hq, hkv, ctx_lens, max_context_len, token_budget, is_long_text=True, d=128,
               compact=False):
    """One layer of Quest on the inputs of oracle/cases.py: prefill bounds via on_kv_stored (two
    chunks that split a page), decode-page completion via on_forward_end, then build_decode_view."""
    from sparsevllm.utils.context import set_context
    from sparsevllm.engine.cache_manager import quest as quest_mod

    inp = cases.make_quest_inputs(seed=seed, bsz=bsz, hq=hq, hkv=hkv, ctx_lens=ctx_lens,
                                  max_context_len=max_context_len, d=d, compact=compact)
    rows = bsz + 1
    m = make_quest_manager(layers=1, hkv=hkv, d=d, num_pages=inp["num_pages"], rows=rows,
                           max_model_len=max_context_len, token_budget=token_budget)
    m.buffer_req_to_page_slots.copy_(inp["page_table"])
    m.buffer_req_to_token_slots.copy_(inp["token_table"])
    req_indices = inp["req_indices"]
    for b in range(bsz):
        row = int(req_indices[b])
        n = int(ctx_lens[b])
        k, v, slots = inp["k_rows"][b], inp["v_rows"][b], inp["slot_rows"][b]
        n_pref = n - 1
        cut = max(1, min(n_pref - 1, (n_pref // 2) | 5)) if n_pref > 2 else n_pref
        set_context(True)
        for lo, hi in ((0, cut), (cut, n_pref)):
            if hi <= lo:
                continue
            sl = slots[lo:hi].to(torch.long)
            m.kv_cache[0, 0].index_copy_(0, sl, k[lo:hi])
            m.kv_cache[1, 0].index_copy_(0, sl, v[lo:hi])
            m.on_kv_stored(0, k[lo:hi], slots[lo:hi])
        # decode step stores the last token; on_forward_end refreshes a page it completes
        set_context(False)
        sl = slots[n - 1:n].to(torch.long)
        m.kv_cache[0, 0].index_copy_(0, sl, k[n - 1:n])
        m.kv_cache[1, 0].index_copy_(0, sl, v[n - 1:n])
        m.row_seq_lens[row] = n
        m.seq_id_to_row[100 + b] = row
    m.buffer_req_to_page_slots_cpu = m.buffer_req_to_page_slots.numpy()
    seqs = [SimpleNamespace(seq_id=100 + b) for b in range(bsz)]
    # on_forward_end -> super().on_forward_end / _poll_prefix_offload are engine plumbing; run the
    # Quest body through the class but neutralise the two parent hooks.
    m._poll_prefix_offload = lambda: None
    orig = None
    for cls in quest_mod.QuestCacheManager.__mro__[1:]:
        if "on_forward_end" in cls.__dict__:
            orig = (cls, cls.__dict__["on_forward_end"])
            break
    if orig is not None:
        setattr(orig[0], "on_forward_end", lambda self, seqs, is_prefill: None)
    try:
        m.on_forward_end(seqs, False)
    finally:
        if orig is not None:
            setattr(orig[0], "on_forward_end", orig[1])
    k_cache, v_cache = cases.quest_fill_cache(inp, hkv, d)
    assert torch.equal(k_cache, m.kv_cache[0, 0]) and torch.equal(v_cache, m.kv_cache[1, 0])

    q = inp["q"]
    context_lens = inp["context_lens"]
    m.layer_batch_state.max_context_len = max_context_len
    set_context(False, is_long_text=is_long_text)
    captured = {}
    orig_score = quest_mod.QuestCacheManager._score_pages_batched

    def spy(q_heads, page_max, page_min, num_kv_heads):
        out = orig_score(q_heads, page_max, page_min, num_kv_heads)
        captured["scores_raw"] = out.clone()
        return out

    quest_mod.QuestCacheManager._score_pages_batched = staticmethod(spy)
    try:
        packed, req_out, lens = m.build_decode_view(
            0, q, m.buffer_req_to_token_slots, req_indices, context_lens, num_heads=hq, num_kv_heads=hkv)
    finally:
        quest_mod.QuestCacheManager._score_pages_batched = staticmethod(orig_score)
    out = {
        "digest": np.array(cases.digest(q, k_cache, v_cache, inp["page_table"], inp["token_table"])),
        "scores_raw": bits(captured["scores_raw"]),
        "packed_slots": packed.numpy().astype(np.int32), "req_out": req_out.numpy().astype(np.int32),
        "lens": lens.numpy().astype(np.int32),
    }
    if compact:
        # BASELINE-size case: the bounds are committed as a digest (max plane, min plane, physical page order;
        # never-completed pages are zero), and the reference's Triton decode kernels (interpreter, fp32-upcast
        # inputs) run on the reference's own packed slots so the attention launch is pinned at this size too
        out["bounds_digest"] = np.array(cases.digest(m.metadata_cache[0, 0], m.metadata_cache[1, 0]))
        from sparsevllm.kernels.triton.flash_decoding_stage2 import flash_decode_stage2
        from sparsevllm.kernels.triton.gqa_flash_decoding_stage1 import flash_decode_stage1
        block_seq, width = 256, int(packed.shape[1])
        nblk = (width + block_seq - 1) // block_seq
        mid_o = torch.zeros(bsz, hq, nblk, d, dtype=torch.float32)
        mid_lse = torch.zeros(bsz, hq, nblk, dtype=torch.float32)
        qf = q.float()
        flash_decode_stage1(qf, k_cache.float(), v_cache.float(), packed, req_out, lens, width, mid_o, mid_lse, block_seq)
        o = torch.empty_like(qf)
        flash_decode_stage2(mid_o, mid_lse, lens, o, block_seq)
        out["attn_out_f32"] = o.numpy()
    else:
        out["bounds_max"] = bits(m.metadata_cache[0, 0])
        out["bounds_min"] = bits(m.metadata_cache[1, 0])
    GOLDEN.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(GOLDEN / f"{name}.npz", **out)
    print(f"[quest] {name}: packed {tuple(packed.shape)} lens {lens.tolist()[:4]}")


def attention_case(name, *, seed, bsz, hq, hkv, lens, width, num_slots, with_score, block_seq=256, d=128):
    """Reference Triton stage1(+2-D score) + stage2 under the interpreter, fp32-upcast inputs."""
    from sparsevllm.kernels.triton.flash_decoding_stage2 import flash_decode_stage2
    from sparsevllm.kernels.triton.gqa_flash_decoding_stage1 import (
        flash_decode_stage1, flash_decode_stage1_with_score)

    inp = cases.make_attn_inputs(seed=seed, bsz=bsz, hq=hq, hkv=hkv, lens=lens, width=width,
                                 num_slots=num_slots, d=d)
    q, k_cache, v_cache = inp["q"], inp["k_cache"], inp["v_cache"]
    table, req, seq_lens = inp["slot_table"], inp["req_indices"], inp["seq_lens"]
    max_len = width
    nblk = (max_len + block_seq - 1) // block_seq
    mid_o = torch.zeros(bsz, hq, nblk, d, dtype=torch.float32)
    mid_lse = torch.zeros(bsz, hq, nblk, dtype=torch.float32)
    qf, kf, vf = q.float(), k_cache.float(), v_cache.float()
    score = None
    if with_score:
        score = torch.full((bsz, width), -1e20, dtype=torch.float32)
        flash_decode_stage1_with_score(qf, kf, vf, table, req, seq_lens, max_len, mid_o, mid_lse, score, block_seq)
    else:
        flash_decode_stage1(qf, kf, vf, table, req, seq_lens, max_len, mid_o, mid_lse, block_seq)
    o = torch.empty_like(qf)
    flash_decode_stage2(mid_o, mid_lse, seq_lens, o, block_seq)
    out = {
        "digest": np.array(cases.digest(q, k_cache, v_cache, table)),
        "block_seq": np.int32(block_seq), "mid_o": mid_o.numpy(), "mid_lse": mid_lse.numpy(),
        "out_f32": o.numpy(),
    }
    if with_score:
        out["attn_score_raw"] = score.numpy().copy()
        # layer-end normalisation exactly as SparseController.on_layer_attention_end does it
        # (engine/sparse_controller.py:761-767)
        norm = score.clone()
        norm.mul_(float(d ** -0.5))
        torch.softmax(norm, dim=-1, out=norm)
        out["attn_score_norm"] = norm.numpy()
    np.savez_compressed(GOLDEN / f"{name}.npz", **out)
    print(f"[attn] {name}: o {tuple(o.shape)} nblk {nblk}")


def h2o_case(name, *, seed, rows, kv_len, budget, ratio, ties):
    from sparsevllm.engine.cache_manager.h2o import H2OCacheManager

    scores = cases.make_h2o_scores(seed=seed, rows=rows, kv_len=kv_len, ties=ties)
    keep_b = H2OCacheManager.select_h2o_indices_batch(scores, budget=budget, recent_ratio=ratio)
    keep_1 = torch.stack([H2OCacheManager.select_h2o_indices(scores[r], budget=budget, recent_ratio=ratio)
                          for r in range(rows)])
    assert torch.equal(keep_b, keep_1)
    np.savez_compressed(GOLDEN / f"{name}.npz", digest=np.array(cases.digest(scores)),
                        keep=keep_b.numpy())
    print(f"[h2o] {name}: keep {tuple(keep_b.shape)}")


def h2o_prefill_case(name, **cfg):
    """Reference Triton prefill scoring (three kernels, probability mode) under the interpreter, fp32-upcast
    inputs, with the arguments H2OCacheManager.collect_prefill_attention_score passes (h2o.py:849-865)."""
    from sparsevllm.kernels.triton.prefill_score import prefill_score_fwd

    inp = cases.make_h2o_prefill_inputs(**cfg)
    bsz = len(cfg["rows"])
    score = torch.empty(bsz, inp["max_context_len"], dtype=torch.float32)
    prefill_score_fwd(inp["q"].float(), inp["k_cache"].float(), score, inp["req_indices"], inp["start_loc"],
                      inp["context_lens"], inp["prompt_cache_lens"], inp["max_query_len"], inp["slot_table"],
                      inp["score_starts"], inp["score_ends"], candidate_start=0, num_recent_tokens=0,
                      score_mode="probability")
    # the dense prefill attention of the same chunk (context_attention_fwd, fp32-upcast inputs under the interpreter)
    from sparsevllm.kernels.triton.context_flashattention_nopad import context_attention_fwd
    qf = inp["q"].float()
    o = torch.zeros_like(qf)
    chunk_lens = inp["context_lens"] - inp["prompt_cache_lens"]
    context_attention_fwd(qf, inp["k_cache"].float(), inp["v_cache"].float(), o, inp["req_indices"], inp["start_loc"],
                          inp["context_lens"], inp["prompt_cache_lens"], int(chunk_lens.max()), inp["slot_table"])
    np.savez_compressed(GOLDEN / f"{name}.npz", digest=np.array(cases.digest(inp["q"], inp["k_cache"], inp["slot_table"])),
                        score=score.numpy(), v_digest=np.array(cases.digest(inp["v_cache"])),
                        attn_out_f16=o[cases.prefill_golden_tokens(inp)].numpy().astype(np.float16))
    print(f"[h2o prefill] {name}: score {tuple(score.shape)} sums {score.sum(dim=1).tolist()} attn {tuple(o.shape)}")


def main():
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for name, cfg in {**cases.QUEST_CASES, **cases.QUEST_BASELINE_CASES}.items():
        quest_case(name, **cfg)
    for name, cfg in cases.ATTN_CASES.items():
        attention_case(name, **cfg)
    for name, cfg in cases.H2O_CASES.items():
        h2o_case(name, **cfg)
    for name, cfg in cases.H2O_PREFILL_CASES.items():
        h2o_prefill_case(name, **cfg)


if __name__ == "__main__":
    main()
