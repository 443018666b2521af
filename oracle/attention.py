"""Oracle: GQA split-K decode attention (stage 1 ± 2-D head-max score, stage 2 combine).

Restates kernels/triton/gqa_flash_decoding_stage1.py:6-121 (plain), 211-295 (2-D score) and
kernels/triton/flash_decoding_stage2.py:8-46 in torch fp32.  Conventions that matter for
parity and are kept: QK in fp32 from bf16 inputs, logits scaled by sm_scale AFTER the raw
score is taken, tile-wise (BLOCK_N tokens) online softmax, softmax numerators cast to the V
dtype before PV (line 97), every split normalised by its own sum (line 113), lse = m + log(sum),
empty split -> (0, -inf).  Test infrastructure only — see oracle/__init__.py.
"""
from __future__ import annotations

import math

import torch


@torch.no_grad()
def gqa_decode_stage1(
    q: torch.Tensor,
    k_cache: torch.Tensor,
    v_cache: torch.Tensor,
    slot_table: torch.Tensor,
    req_indices: torch.Tensor,
    seq_lens: torch.Tensor,
    *,
    num_blocks: int,
    block_seq: int = 256,
    block_n: int = 16,
    sm_scale: float | None = None,
    attn_score: torch.Tensor | None = None,
) -> tuple[torch.Tensor, torch.Tensor]:
    """Returns (mid_o [B,Hq,num_blocks,D] fp32, mid_lse [B,Hq,num_blocks] fp32).

    ``attn_score`` (fp32 [B, W], pre-filled by the caller, reference: -1e20) is max-updated
    in place with the raw, unscaled max over ALL local q heads of q.k[n]
    (gqa_flash_decoding_stage1.py:262-270: per-kv-head max then ``tl.atomic_max``).
    """
    bsz, hq, d = q.shape
    hkv = k_cache.shape[1]
    g = hq // hkv
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(d)  # wrapper line 353/410: 1.0 / (Lk ** 0.5)
    p_dtype = v_cache.dtype
    mid_o = torch.zeros((bsz, hq, num_blocks, d), dtype=torch.float32)
    mid_lse = torch.full((bsz, hq, num_blocks), -float("inf"), dtype=torch.float32)
    for b in range(bsz):
        n = int(seq_lens[b])
        if n <= 0:
            continue
        slots = slot_table[int(req_indices[b]), :n].to(torch.long)
        k = k_cache.index_select(0, slots).float()  # [n, Hkv, D]
        v = v_cache.index_select(0, slots).float()
        qf = q[b].float().view(hkv, g, d)
        raw = torch.einsum("hgd,nhd->hgn", qf, k)  # fp32 accumulate of bf16 products
        if attn_score is not None:
            torch.maximum(attn_score[b, :n], raw.amax(dim=(0, 1)), out=attn_score[b, :n])
        s = raw * sm_scale
        for blk in range(num_blocks):
            s0 = blk * block_seq
            s1 = min(n, s0 + block_seq)
            if s1 <= s0:
                continue  # neutral partial already in place (lines 107-120)
            m = torch.full((hkv, g), -float("inf"))
            l = torch.zeros((hkv, g))
            acc = torch.zeros((hkv, g, d))
            for t0 in range(s0, s1, block_n):
                t1 = min(s1, t0 + block_n)
                st = s[:, :, t0:t1]
                m_new = torch.maximum(m, st.amax(dim=-1))
                p = torch.exp(st - m_new[..., None])
                alpha = torch.exp(m - m_new)
                acc = acc * alpha[..., None] + torch.einsum(
                    "hgn,nhd->hgd", p.to(p_dtype).float(), v[t0:t1]
                )
                l = l * alpha + p.sum(dim=-1)
                m = m_new
            mid_o[b, :, blk] = (acc / l[..., None]).reshape(hq, d)
            mid_lse[b, :, blk] = (m + torch.log(l)).reshape(hq)
    return mid_o, mid_lse


@torch.no_grad()
def decode_stage2(mid_o: torch.Tensor, mid_lse: torch.Tensor, seq_lens: torch.Tensor, *, block_seq: int,
                  out_dtype: torch.dtype = torch.bfloat16) -> torch.Tensor:
    """flash_decoding_stage2.py:8-46: sequential LSE-weighted merge over ceil(len/block_seq) blocks."""
    bsz, hq, _, d = mid_o.shape
    out = torch.zeros((bsz, hq, d), dtype=torch.float32)
    for b in range(bsz):
        n = int(seq_lens[b])
        nblk = 0 if n <= 0 else (n + block_seq - 1) // block_seq
        m = torch.full((hq,), -float("inf"))
        l = torch.zeros((hq,))
        acc = torch.zeros((hq, d))
        for i in range(nblk):
            lse = mid_lse[b, :, i]
            m_new = torch.maximum(lse, m)
            old = torch.exp(m - m_new)
            w = torch.exp(lse - m_new)
            acc = acc * old[:, None] + w[:, None] * mid_o[b, :, i]
            l = l * old + w
            m = m_new
        out[b] = acc / l[:, None]
    return out.to(out_dtype)


@torch.no_grad()
def gqa_decode(q, k_cache, v_cache, slot_table, req_indices, seq_lens, *, max_len_in_batch: int,
               block_seq: int = 256, block_n: int = 16, sm_scale: float | None = None,
               attn_score: torch.Tensor | None = None) -> torch.Tensor:
    """stage1 + stage2 as TritonAttentionBackend.run_decode chains them
    (layers/attention_backend.py:283-348; workspace blocks per layers/attention.py:216-236)."""
    num_blocks = (int(max_len_in_batch) + block_seq - 1) // block_seq
    mid_o, mid_lse = gqa_decode_stage1(
        q, k_cache, v_cache, slot_table, req_indices, seq_lens, num_blocks=num_blocks,
        block_seq=block_seq, block_n=block_n, sm_scale=sm_scale, attn_score=attn_score)
    return decode_stage2(mid_o, mid_lse, seq_lens, block_seq=block_seq, out_dtype=q.dtype)


@torch.no_grad()
def dense_softmax_attention_f64(q, k_cache, v_cache, slot_table, req_indices, seq_lens, sm_scale=None):
    """Plain float64 softmax attention over the same slots (no tiling, no bf16 P).

    Not reference code: an independent cross-check that both the restatement above and the
    CUDA kernel stay within the stated tolerance of the mathematically exact answer.
    """
    bsz, hq, d = q.shape
    hkv = k_cache.shape[1]
    g = hq // hkv
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(d)
    out = torch.zeros((bsz, hq, d), dtype=torch.float64)
    for b in range(bsz):
        n = int(seq_lens[b])
        if n <= 0:
            continue
        slots = slot_table[int(req_indices[b]), :n].to(torch.long)
        k = k_cache.index_select(0, slots).double()
        v = v_cache.index_select(0, slots).double()
        s = torch.einsum("hgd,nhd->hgn", q[b].double().view(hkv, g, d), k) * sm_scale
        p = torch.softmax(s, dim=-1)
        out[b] = torch.einsum("hgn,nhd->hgd", p, v).reshape(hq, d)
    return out


def prefill_attention(q, k_cache, v_cache, slot_table, req_indices, start_loc, context_lens, prompt_cache_lens):
    """Dense causal prefill attention of a chunk over its paged row: ``context_attention_fwd``
    (kernels/triton/context_flashattention_nopad.py:11-79, 242-302).  Query m of row b sits at physical position
    m + prompt_cache_len and attends to kv positions <= its own through the slot table; logits q.k * D^-0.5,
    softmax in fp32, numerators cast to V's dtype before PV (line 70), output in q's dtype.  q [T, Hq, D]."""
    hq, d = int(q.shape[1]), int(q.shape[2])
    hkv = int(k_cache.shape[1])
    g = hq // hkv
    out = torch.zeros_like(q)
    scale = float(d) ** -0.5
    for b in range(int(req_indices.numel())):
        ctx, cache_len = int(context_lens[b]), int(prompt_cache_lens[b])
        n = ctx - cache_len
        if n <= 0:
            continue
        slots = slot_table[int(req_indices[b]), :ctx].long()
        k = k_cache[slots].float().repeat_interleave(g, dim=1)      # [ctx, Hq, D]
        v = v_cache[slots].repeat_interleave(g, dim=1)
        s0 = int(start_loc[b])
        qv = q[s0: s0 + n].float()
        logits = torch.einsum("mhd,nhd->hmn", qv, k) * scale
        causal = (torch.arange(n)[:, None] + cache_len) >= torch.arange(ctx)[None, :]
        logits = logits.masked_fill(~causal[None], float("-inf"))
        m = logits.max(dim=-1, keepdim=True).values
        p = torch.exp(logits - m)
        l = p.sum(dim=-1, keepdim=True)
        o = torch.einsum("hmn,nhd->mhd", p.to(v.dtype).float(), v.float()) / l.permute(1, 0, 2)
        out[s0: s0 + n] = o.to(q.dtype)
    return out
