"""GPU parity: sparse paged GQA decode attention (split-K + fused combine, optional H2O 2-D score)
and the H2O normalise / accumulate / select / compact kernels, through the C ABI.
Float tolerance (north_star): 2e-2 abs / 1e-2 rel on bf16 outputs."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from oracle import attention as oattn
from oracle import cases, h2o as oh2o

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
ATOL, RTOL = 2e-2, 1e-2


def _run_attn(inp, *, max_ctas, hpc, with_score, width):
    from sparse_vllm_b200 import ops
    q = inp["q"].to(DEV)
    bsz, hq, _ = q.shape
    hkv = inp["k_cache"].shape[1]
    out = torch.full_like(q, float("nan"))
    ws = ops.attn_workspace(bsz, hq, hkv, DEV, max_ctas=max_ctas)
    ws[(bsz * hkv * 4 + 255) // 256 * 64:] = 0x7FC00000  # partial records start out as NaN
    score = torch.full((bsz, width), -1e20, dtype=torch.float32, device=DEV) if with_score else None
    for _ in range(2):  # second run must see zeroed tickets and give the same answer
        if score is not None:
            score.fill_(-1e20)
        ops.sparse_decode_attn(q, inp["k_cache"].to(DEV), inp["v_cache"].to(DEV), inp["slot_table"].to(DEV),
                               inp["req_indices"].to(DEV), inp["seq_lens"].to(DEV), out=out, workspace=ws,
                               max_len=width, attn_score=score, heads_per_cta=hpc)
    torch.cuda.synchronize()
    assert int(ws[: bsz * hkv].abs().sum()) == 0
    return out.cpu(), None if score is None else score.cpu()


def _variants(hkv, width):
    """(heads_per_cta, CTA count): 1 CTA = every row complete, a few CTAs = rows cut mid-way, None = one per SM."""
    out = []
    for hpc in (0, 1, 2, 4, 8):
        if hpc and hkv % hpc:
            continue
        for max_ctas in (None, 1, 3, 37):
            out.append((hpc, max_ctas))
    return out


@pytest.mark.parametrize("name", list(cases.ATTN_CASES))
def test_attention_matches_reference(name, golden):
    cfg = cases.ATTN_CASES[name]
    inp = cases.make_attn_inputs(**cfg)
    g = golden(name)
    assert str(g["digest"]) == cases.digest(inp["q"], inp["k_cache"], inp["v_cache"], inp["slot_table"])
    ref = torch.from_numpy(g["out_f32"])
    live = [b for b, n in enumerate(cfg["lens"]) if n > 0]
    dead = [b for b, n in enumerate(cfg["lens"]) if n == 0]
    oracle_out = oattn.gqa_decode(inp["q"], inp["k_cache"], inp["v_cache"], inp["slot_table"], inp["req_indices"],
                                  inp["seq_lens"], max_len_in_batch=cfg["width"]).float()
    for hpc, max_ctas in _variants(cfg["hkv"], cfg["width"]):
        out, score = _run_attn(inp, max_ctas=max_ctas, hpc=hpc, with_score=cfg["with_score"], width=cfg["width"])
        tag = f"{name} hpc={hpc} ctas={max_ctas}"
        torch.testing.assert_close(out[live].float(), ref[live], atol=ATOL, rtol=RTOL, msg=lambda m: f"{tag}: {m}")
        torch.testing.assert_close(out[live].float(), oracle_out[live], atol=ATOL, rtol=RTOL, msg=lambda m: f"{tag}: {m}")
        if dead:  # an empty row yields zeros here (the reference divides 0/0); it must not be NaN
            assert bool((out[dead] == 0).all()), tag
        if cfg["with_score"]:
            ref_score = torch.from_numpy(g["attn_score_raw"])
            # raw q.k maxima: fp32 accumulation-order noise only
            torch.testing.assert_close(score, ref_score, atol=2e-3, rtol=1e-4, msg=lambda m: f"{tag} score: {m}")
            for b, n in enumerate(cfg["lens"]):
                assert bool((score[b, n:] == -1e20).all()), f"{tag}: score written beyond the row length"


def test_attention_masked_slots_and_strided_io():
    """Entries < 0 in the slot table are masked tokens; q/out may be strided in the batch dimension."""
    from sparse_vllm_b200 import ops
    cfg = dict(seed=77, bsz=2, hq=16, hkv=2, lens=[200, 150], width=256, num_slots=512)
    inp = cases.make_attn_inputs(**cfg)
    table = inp["slot_table"].clone()
    masked = {1: [3, 17, 40, 199], 2: [0, 149]}
    for row, cols in masked.items():
        table[row, cols] = -1
    # oracle on the compacted rows (masked tokens removed)
    t2 = torch.zeros_like(table)
    lens2 = []
    for b in range(2):
        row = int(inp["req_indices"][b])
        keep = [c for c in range(cfg["lens"][b]) if int(table[row, c]) >= 0]
        t2[row, :len(keep)] = table[row, keep]
        lens2.append(len(keep))
    ref = oattn.gqa_decode(inp["q"], inp["k_cache"], inp["v_cache"], t2, inp["req_indices"],
                           torch.tensor(lens2, dtype=torch.int32), max_len_in_batch=256).float()
    q_big = torch.zeros(2, 3, 16, 128, dtype=torch.bfloat16, device=DEV)
    q_big[:, 1] = inp["q"].to(DEV)
    out_big = torch.zeros_like(q_big)
    for max_ctas in (None, 5):
        out_big.zero_()
        ops.sparse_decode_attn(q_big[:, 1], inp["k_cache"].to(DEV), inp["v_cache"].to(DEV), table.to(DEV),
                               inp["req_indices"].to(DEV), inp["seq_lens"].to(DEV), out=out_big[:, 1],
                               workspace=ops.attn_workspace(2, 16, 2, DEV, max_ctas=max_ctas))
        torch.cuda.synchronize()
        torch.testing.assert_close(out_big[:, 1].cpu().float(), ref, atol=ATOL, rtol=RTOL)
    assert not out_big[:, 0].any() and not out_big[:, 2].any()


@pytest.mark.parametrize("hq,hkv,bsz,tokens", [(28, 4, 8, 4096), (32, 8, 4, 2176), (64, 8, 2, 4096), (8, 1, 4, 4096)])
def test_attention_full_width_properties(hq, hkv, bsz, tokens):
    """BASELINE selection widths: every CTA-count/head-group decomposition gives the same answer, and it
    matches fp32 torch softmax attention over the same slots."""
    from sparse_vllm_b200 import ops
    gen = torch.Generator(device=DEV).manual_seed(99)
    num_slots = bsz * tokens + 64
    k = torch.randn(num_slots, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
    v = torch.randn(num_slots, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
    q = torch.randn(bsz, hq, 128, device=DEV, generator=gen).to(torch.bfloat16)
    table = torch.randperm(num_slots, device=DEV, generator=gen)[: bsz * tokens].to(torch.int32).view(bsz, tokens)
    lens = torch.tensor([tokens - 5 * b for b in range(bsz)], dtype=torch.int32, device=DEV)
    req = torch.arange(bsz, dtype=torch.int32, device=DEV)
    g = hq // hkv
    ref = torch.empty(bsz, hq, 128, device=DEV)
    for b in range(bsz):
        n = int(lens[b])
        kk = k[table[b, :n].long()].float()
        vv = v[table[b, :n].long()].float()
        s = torch.einsum("hgd,nhd->hgn", q[b].float().view(hkv, g, 128), kk) * 128 ** -0.5
        ref[b] = torch.einsum("hgn,nhd->hgd", torch.softmax(s, -1), vv).reshape(hq, 128)
    outs = []
    for hpc in (0, 1, hkv):
        for max_ctas in (None, 64, bsz):
            out = torch.empty_like(q)
            ws = ops.attn_workspace(bsz, hq, hkv, DEV, max_ctas=max_ctas)
            ops.sparse_decode_attn(q, k, v, table, req, lens, out=out, workspace=ws, heads_per_cta=hpc)
            torch.cuda.synchronize()
            torch.testing.assert_close(out.float(), ref, atol=ATOL, rtol=RTOL)
            outs.append(out.float())
    for o in outs[1:]:
        torch.testing.assert_close(o, outs[0], atol=ATOL, rtol=RTOL)


# ------------------------------------------------------------------------------------- H2O
def test_h2o_normalize_matches_reference(golden):
    from sparse_vllm_b200 import ops
    for name in ("attn_g4_score", "attn_g8_empty_row"):
        g = golden(name)
        raw = torch.from_numpy(g["attn_score_raw"]).to(DEV)
        out = ops.h2o_normalize_accumulate(raw, 128 ** -0.5)
        torch.cuda.synchronize()
        assert out.data_ptr() == raw.data_ptr()
        torch.testing.assert_close(raw.cpu(), torch.from_numpy(g["attn_score_norm"]), atol=1e-6, rtol=1e-4)


@pytest.mark.parametrize("width", [2176, 320, 40000])
def test_h2o_normalize_accumulate_fused(width):
    from sparse_vllm_b200 import ops
    gen = torch.Generator().manual_seed(width)
    bsz = 5
    kv_lens = torch.tensor([min(width, 50 + 37 * b) for b in range(bsz)], dtype=torch.int32)
    raw = torch.full((bsz, width), oh2o.SCORE_FILL)
    for b in range(bsz):
        raw[b, : int(kv_lens[b])] = torch.randn(int(kv_lens[b]), generator=gen) * 30
    prev = torch.rand(bsz, width, generator=gen)
    exp_norm = oh2o.normalize_scores_(raw.clone(), 128 ** -0.5)
    exp_cum = prev.clone()
    for b in range(bsz):
        n = int(kv_lens[b])
        exp_cum[b, :n] = oh2o.accumulate_all_layers(exp_norm[None, b:b + 1], prev[None, b:b + 1, : n - 1], n)[0, 0]
    score, cum = raw.to(DEV), prev.to(DEV)
    ops.h2o_normalize_accumulate(score, 128 ** -0.5, cum, kv_lens.to(DEV))
    torch.cuda.synchronize()
    torch.testing.assert_close(score.cpu(), exp_norm, atol=1e-6, rtol=1e-4)
    torch.testing.assert_close(cum.cpu(), exp_cum, atol=1e-6, rtol=1e-4)


def _select_compact(scores, slots, budget, recent):
    from sparse_vllm_b200 import ops
    rows, kv_len = scores.shape
    cum = scores.clone().to(DEV)
    tab = slots.clone().to(DEV)
    offs = (torch.arange(rows, dtype=torch.int64) * kv_len).to(DEV)
    drop = kv_len - budget
    free = torch.full((rows * drop + 4,), -1, dtype=torch.int32, device=DEV)
    free_rows = (torch.arange(rows, dtype=torch.int64) * drop + 2).to(DEV)
    keep = torch.empty(rows, budget, dtype=torch.int32, device=DEV)
    ops.h2o_select_compact(cum, offs, tab, offs, kv_len=kv_len, budget=budget, recent=recent,
                           free_stack=free, free_rows=free_rows, keep_out=keep)
    torch.cuda.synchronize()
    return cum.cpu(), tab.cpu(), free.cpu(), keep.cpu()


@pytest.mark.parametrize("name", list(cases.H2O_CASES))
def test_h2o_select_compact_matches_reference(name, golden):
    cfg = cases.H2O_CASES[name]
    scores = cases.make_h2o_scores(**cfg)
    g = golden(name)
    assert str(g["digest"]) == cases.digest(scores)
    rows, kv_len = scores.shape
    budget = cfg["budget"]
    recent = oh2o.recent_count(budget, cfg["ratio"], kv_len)
    slots = (torch.arange(rows * kv_len, dtype=torch.int32).view(rows, kv_len) * 7 + 3) % 100003
    cum, tab, free, keep = _select_compact(scores, slots, budget, recent)
    ref_keep = torch.from_numpy(g["keep"])
    assert torch.equal(keep.long(), ref_keep), "kept indices differ from select_h2o_indices_batch (stable order)"
    exp_slots, exp_scores, exp_drop = oh2o.compact_rows(slots, scores, ref_keep)
    assert torch.equal(tab, exp_slots)
    assert torch.equal(cum[:, :budget], exp_scores) and not cum[:, budget:].any()
    drop = kv_len - budget
    assert torch.equal(free[2: 2 + rows * drop].view(rows, drop), exp_drop)
    assert free[:2].tolist() == [-1, -1] and free[-2:].tolist() == [-1, -1]


def test_h2o_select_compact_full_size():
    """cfg 4 geometry: 32 layers x 64 rows, kv_len 2176 -> budget 2048 (1024 heavy + 1024 recent)."""
    gen = torch.Generator().manual_seed(4)
    rows, kv_len, budget = 256, 2176, 2048
    scores = torch.rand(rows, kv_len, generator=gen)
    scores[:, ::7] = 0.25  # ties
    slots = torch.stack([torch.randperm(1 << 20, generator=gen)[:kv_len] for _ in range(rows)]).to(torch.int32)
    recent = oh2o.recent_count(budget, 0.5, kv_len)
    cum, tab, free, keep = _select_compact(scores, slots, budget, recent)
    ref_keep = oh2o.select_indices_batch(scores, budget=budget, recent_ratio=0.5)
    assert torch.equal(keep.long(), ref_keep)
    exp_slots, exp_scores, exp_drop = oh2o.compact_rows(slots, scores, ref_keep)
    assert torch.equal(tab, exp_slots) and torch.equal(cum[:, :budget], exp_scores)
    assert torch.equal(free[2: 2 + rows * (kv_len - budget)].view(rows, -1), exp_drop)


@pytest.mark.parametrize("name", list(cases.H2O_PREFILL_CASES))
def test_h2o_prefill_score_matches_reference(name, golden):
    """svb200_h2o_prefill_score against the reference's Triton prefill scoring (golden: interpreter, fp32-upcast
    inputs) and the oracle, plus the fused cumulative update (expand with zeros, weight = scoring queries)."""
    from sparse_vllm_b200 import ops
    cfg = cases.H2O_PREFILL_CASES[name]
    inp = cases.make_h2o_prefill_inputs(**cfg)
    g = golden(name)
    assert str(g["digest"]) == cases.digest(inp["q"], inp["k_cache"], inp["slot_table"])
    bsz, width = len(cfg["rows"]), inp["max_context_len"]
    dev = {k: v.to(DEV) for k, v in inp.items() if isinstance(v, torch.Tensor)}
    score = torch.full((bsz, width + 3), float("nan"), dtype=torch.float32, device=DEV)
    gen = torch.Generator().manual_seed(1)
    prev = torch.rand(bsz + 1, width + 5, generator=gen)  # cumulative rows (table rows), garbage beyond the prefix
    cum = prev.clone().to(DEV)
    for _ in range(2):  # idempotent w.r.t. the score buffer (the launcher zero-fills it)
        ops.h2o_prefill_score(dev["q"], dev["k_cache"], dev["slot_table"], dev["req_indices"], dev["start_loc"],
                              dev["context_lens"], dev["prompt_cache_lens"], dev["score_starts"], dev["score_ends"], score,
                              max_query_len=inp["max_query_len"], max_context_len=width)
    ops.h2o_prefill_score(dev["q"], dev["k_cache"], dev["slot_table"], dev["req_indices"], dev["start_loc"],
                          dev["context_lens"], dev["prompt_cache_lens"], dev["score_starts"], dev["score_ends"], score,
                          max_query_len=inp["max_query_len"], max_context_len=width, cumulative=cum, cum_rows=dev["req_indices"])
    torch.cuda.synchronize()
    ref = torch.from_numpy(g["score"])
    orc = oh2o.prefill_window_score(inp["q"], inp["k_cache"], inp["slot_table"], inp["req_indices"], inp["start_loc"],
                                    inp["context_lens"], inp["prompt_cache_lens"], inp["score_starts"], inp["score_ends"])
    got = score.cpu()
    for b, n in enumerate(inp["context_lens"].tolist()):
        torch.testing.assert_close(got[b, :n], ref[b, :n], atol=2e-5, rtol=2e-3)
        torch.testing.assert_close(got[b, :n], orc[b, :n], atol=2e-5, rtol=2e-3)
        row = int(inp["req_indices"][b])
        cache_len = int(inp["prompt_cache_lens"][b])
        weight = float(inp["score_ends"][b] - inp["score_starts"][b])
        want = oh2o.accumulate_prefill(prev[row, :cache_len] if cache_len else None, got[b], new_len=n, weight=weight)
        torch.testing.assert_close(cum[row, :n].cpu(), want, atol=1e-5, rtol=1e-5)
        assert torch.equal(cum[row, n:].cpu(), prev[row, n:]), "cumulative written beyond the row"


@pytest.mark.parametrize("name", list(cases.H2O_PREFILL_CASES))
def test_prefill_attention_matches_reference(name, golden):
    """svb200_prefill_attn (dense causal chunk attention over the paged row) against the reference's Triton
    context_attention_fwd (golden) and the oracle on every token."""
    from sparse_vllm_b200 import ops
    inp = cases.make_h2o_prefill_inputs(**cases.H2O_PREFILL_CASES[name])
    g = golden(name)
    assert str(g["v_digest"]) == cases.digest(inp["v_cache"])
    dev = {k: v.to(DEV) for k, v in inp.items() if isinstance(v, torch.Tensor)}
    chunk = inp["context_lens"] - inp["prompt_cache_lens"]
    out = torch.full_like(dev["q"], float("nan"))
    ops.prefill_attn(dev["q"], dev["k_cache"], dev["v_cache"], dev["slot_table"], dev["req_indices"], dev["start_loc"],
                     dev["context_lens"], dev["prompt_cache_lens"], max_chunk_len=int(chunk.max()), out=out)
    torch.cuda.synchronize()
    ref = oattn.prefill_attention(inp["q"], inp["k_cache"], inp["v_cache"], inp["slot_table"], inp["req_indices"],
                                  inp["start_loc"], inp["context_lens"], inp["prompt_cache_lens"])
    torch.testing.assert_close(out.cpu().float(), ref.float(), atol=ATOL, rtol=RTOL)
    tok = cases.prefill_golden_tokens(inp)
    torch.testing.assert_close(out.cpu()[tok].float(), torch.from_numpy(g["attn_out_f16"]).float(), atol=ATOL, rtol=RTOL)
