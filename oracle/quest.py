"""Oracle: Quest page bounds, page scoring, top-k page selection and slot packing.

Restates engine/cache_manager/quest.py of the reference (line numbers in each docstring).
Test infrastructure only — see oracle/__init__.py.
"""
from __future__ import annotations

import numpy as np
import torch

PAGE = 16


def page_minmax(k_pages: torch.Tensor) -> tuple[torch.Tensor, torch.Tensor]:
    """Per-page (max, min) over the token axis of ``k_pages [n_pages, page, Hkv, D]``.

    quest.py:1628-1644 / 1662-1685 / 1759-1771 all reduce a full page with
    ``torch.aminmax(full_page_k, dim=1)``; partial trailing pages have no bounds.
    """
    mn, mx = torch.aminmax(k_pages, dim=1)
    return mx, mn


def build_bounds_from_cache(k_cache: torch.Tensor, page_size: int = PAGE) -> tuple[torch.Tensor, torch.Tensor]:
    """Bounds of EVERY physical page of one layer's K cache ``[slots, Hkv, D]``.

    What ``metadata_cache[0/1, layer]`` holds (quest.py:96-105) once all pages are complete.
    """
    slots, hkv, d = k_cache.shape
    return page_minmax(k_cache.view(slots // page_size, page_size, hkv, d))


def score_pages(q: torch.Tensor, page_max: torch.Tensor, page_min: torch.Tensor, num_kv_heads: int) -> torch.Tensor:
    """``_score_pages_batched`` (quest.py:1773-1802).

    q [B, Hq, D]; page_max/min [B, Hkv, P, D] (dtype of the bounds, bf16 in production).
    Two batched matmuls whose OUTPUTS are in the bounds dtype (so each is rounded to bf16),
    an in-place add in that dtype, then max over the GQA group and over KV heads -> [B, P].
    """
    b, hq, d = q.shape
    dt = page_max.dtype
    p = page_max.shape[2]
    if hq == num_kv_heads:  # MHA branch, quest.py:1782-1792
        qh = q.to(dt)
        qp = qh.clamp_min(0).reshape(b * hq, 1, d)
        qn = qh.clamp_max(0).reshape(b * hq, 1, d)
        s = torch.bmm(qp, page_max.reshape(b * hq, p, d).transpose(1, 2)).squeeze(1)
        s += torch.bmm(qn, page_min.reshape(b * hq, p, d).transpose(1, 2)).squeeze(1)
        return s.view(b, hq, p).amax(dim=1)
    g = hq // num_kv_heads
    qg = q.view(b, num_kv_heads, g, d).to(dt)
    qp = qg.clamp_min(0).reshape(b * num_kv_heads, g, d)
    qn = qg.clamp_max(0).reshape(b * num_kv_heads, g, d)
    s = torch.bmm(qp, page_max.reshape(b * num_kv_heads, p, d).transpose(1, 2))
    s += torch.bmm(qn, page_min.reshape(b * num_kv_heads, p, d).transpose(1, 2))
    return s.view(b, num_kv_heads, g, p).amax(dim=2).amax(dim=1)


def score_pages_f64(q: torch.Tensor, page_max: torch.Tensor, page_min: torch.Tensor, num_kv_heads: int):
    """Same quantity with float64 dot products: returns (pos, neg) per (b, head, page).

    Not part of the reference; used by the parity checker to tell an exact tie / a 1-ulp
    accumulation-order flip from a real error (SURVEY.md §7 "hard parts").
    """
    b, hq, d = q.shape
    g = hq // num_kv_heads
    qd = q.double().view(b, num_kv_heads, g, d)
    pos = torch.einsum("bhgd,bhpd->bhgp", qd.clamp_min(0), page_max.double())
    neg = torch.einsum("bhgd,bhpd->bhgp", qd.clamp_max(0), page_min.double())
    return pos, neg


def build_decode_view_static(
    q: torch.Tensor,
    bounds_max: torch.Tensor,
    bounds_min: torch.Tensor,
    page_table: torch.Tensor,
    token_table: torch.Tensor | None,
    req_indices: torch.Tensor,
    context_lens: torch.Tensor,
    *,
    token_budget: int,
    num_kv_heads: int,
    max_context_len: int,
    max_pages_per_row: int,
    is_long_text: bool = True,
    page_size: int = PAGE,
):
    """``_build_decode_view_static`` (quest.py:1833-1913) for one layer.

    bounds_max/min [P_total, Hkv, D] indexed by PHYSICAL page slot; page_table
    [rows, max_pages] int32 (-1 empty); token_table [rows, max_model_len] int32 (dense slots,
    only read when not long-text).  Returns a dict with scores (masked, bounds dtype),
    selected logical pages [B, prev+1], packed_slots [B, W] int32, lens [B] int32, or
    ``None`` when the reference returns the dense view unchanged (1850-1861).
    """
    page_budget_base = max(3, int(token_budget) // page_size)
    max_keep = max(int(token_budget), page_budget_base * page_size, page_size)
    if max_context_len <= max_keep:
        return None
    bsz = q.shape[0]
    max_pages = min(max_pages_per_row, (max_context_len + page_size - 1) // page_size)
    prev_budget = min(page_budget_base - 1, max_pages - 1)
    if prev_budget <= 0:
        return None
    num_pages = torch.div(context_lens + page_size - 1, page_size, rounding_mode="floor")
    rows = req_indices.to(torch.long)
    row_page_slots = page_table.index_select(0, rows)[:, :max_pages]
    prev_slots = row_page_slots[:, : max_pages - 1].to(torch.long).clamp_min(0)
    d = bounds_max.shape[-1]
    pmax = bounds_max.index_select(0, prev_slots.reshape(-1)).view(bsz, max_pages - 1, num_kv_heads, d).permute(0, 2, 1, 3)
    pmin = bounds_min.index_select(0, prev_slots.reshape(-1)).view(bsz, max_pages - 1, num_kv_heads, d).permute(0, 2, 1, 3)
    scores = score_pages(q, pmax, pmin, num_kv_heads)
    safe_np = num_pages.to(torch.long).clamp_min(1)
    valid = torch.arange(max_pages - 1)[None, :] < (safe_np - 1)[:, None]
    scores = scores.masked_fill(~valid, -float("inf"))
    top_prev = scores.topk(prev_budget, dim=-1, sorted=False).indices
    last_page = (safe_np - 1)[:, None]
    selected = torch.cat((top_prev, last_page), dim=1)
    sel_slots = row_page_slots.gather(1, selected)
    offs = torch.arange(page_size, dtype=torch.int32)
    sparse_slots = (sel_slots[:, :, None] * page_size + offs[None, None, :]).reshape(bsz, -1)
    last_len = context_lens - (num_pages - 1) * page_size
    sparse_lens = (prev_budget * page_size + last_len).to(torch.int32)
    if is_long_text:
        packed, lens = sparse_slots.to(torch.int32), sparse_lens
    else:
        dense = token_table.index_select(0, rows)[:, :max_keep]
        dense_mask = (context_lens <= int(token_budget)) | (num_pages <= page_budget_base)
        keep = sparse_slots.shape[1]
        packed = torch.empty((bsz, max_keep), dtype=torch.int32)
        packed[:, :keep] = torch.where(dense_mask[:, None], dense[:, :keep], sparse_slots.to(torch.int32))
        if max_keep > keep:
            packed[:, keep:] = dense[:, keep:]
        lens = torch.where(dense_mask, context_lens, sparse_lens).to(torch.int32)
    return {
        "scores": scores,
        "selected_pages": selected,
        "packed_slots": packed,
        "lens": lens,
        "prev_budget": prev_budget,
        "max_pages": max_pages,
        "req_indices": torch.arange(bsz, dtype=torch.int32),
    }


# ------------------------------------------------------------------ checker
def bf16_ulp_distance(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """Distance in bf16 ulps between two bf16 tensors (monotone integer mapping)."""
    def key(x):
        bits = x.contiguous().view(torch.int16).to(torch.int32) & 0xFFFF
        return torch.where(bits >= 0x8000, 0x8000 - bits, bits)
    return (key(a) - key(b)).abs()


def classify_selection(scores_ref: torch.Tensor, sel_ref, sel_got, n_valid: int) -> dict:
    """Compare two top-k page sets of ONE row given the reference (masked) scores.

    Returns counts: ``exact`` (same set), ``tie`` (differences only among pages whose
    reference score equals the k-th score: allowed by north_star), ``bad`` (anything else).
    """
    ref = set(int(x) for x in sel_ref)
    got = set(int(x) for x in sel_got)
    if ref == got:
        return {"exact": 1, "tie": 0, "bad": 0, "detail": ""}
    s = scores_ref.float()
    kth = min(float(s[i]) for i in ref)
    only_ref, only_got = ref - got, got - ref
    ok = all(float(s[i]) == kth for i in only_ref) and all(
        0 <= i < max(n_valid, 1) and float(s[i]) == kth for i in only_got
    )
    detail = f"only_ref={sorted(only_ref)[:8]} only_got={sorted(only_got)[:8]} kth={kth}"
    return {"exact": 0, "tie": int(ok), "bad": int(not ok), "detail": detail}


def classify_selection_ulp(scores_ref: torch.Tensor, sel_ref, sel_got, n_valid: int, ulp_slack: int = 1) -> dict:
    """``classify_selection`` plus the third class of SURVEY.md §7: ``near`` = the two page sets differ only
    in pages whose REFERENCE score sits within ``ulp_slack`` bf16 ulps of the reference's k-th score (a kernel
    that selects on its own scores may see those one rounding step away: cuBLAS / oneDNN / this kernel
    accumulate fp32 in different orders).  Anything else is ``bad`` — in particular a wrong selection made on
    correct scores.  Returns counts exact / tie / near / bad (exactly one is 1)."""
    c = classify_selection(scores_ref, sel_ref, sel_got, n_valid)
    out = {"exact": c["exact"], "tie": c["tie"], "near": 0, "bad": c["bad"], "detail": c["detail"]}
    if not c["bad"] or ulp_slack <= 0:
        return out
    ref, got = set(int(x) for x in sel_ref), set(int(x) for x in sel_got)
    if any(not 0 <= i < max(n_valid, 1) for i in got):
        return out
    kth = scores_ref[sorted(ref)].float().min().to(scores_ref.dtype)
    swapped = torch.tensor(sorted(ref ^ got), dtype=torch.long)
    d = bf16_ulp_distance(scores_ref[swapped], kth.expand(len(swapped)).contiguous())
    if int(d.max()) <= ulp_slack:
        out.update(near=1, bad=0)
    return out


def topk_numpy_lowest_index(scores_row: np.ndarray, k: int) -> np.ndarray:
    """Deterministic top-k used for size-independent checks: ties -> lowest index."""
    order = np.lexsort((np.arange(scores_row.size), -scores_row.astype(np.float64)))
    return np.sort(order[:k])
