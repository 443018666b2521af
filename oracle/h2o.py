"""Oracle: H2O layer-end normalisation, cumulative scores, heavy-hitter selection, compaction.

Restates engine/sparse_controller.py:761-767, engine/cache_manager/h2o.py:479-563,958-1038,
1558-1625 and engine/cache_manager/snapkv.py:1681-1803.  Test infrastructure only.
"""
from __future__ import annotations

import torch

SCORE_FILL = -1e20  # sparse_controller.py:427-497: the [L,B,W] buffer is filled with -1e20


def normalize_scores_(attn_score: torch.Tensor, softmax_scale: float) -> torch.Tensor:
    """sparse_controller.py:766-767: ``attn_score.mul_(scale); softmax(dim=-1, out=attn_score)``."""
    attn_score.mul_(float(softmax_scale))
    torch.softmax(attn_score, dim=-1, out=attn_score)
    return attn_score


def accumulate_all_layers(step_scores: torch.Tensor, previous: torch.Tensor, kv_len: int) -> torch.Tensor:
    """h2o.py:1024-1031: cumulative[..., :kv_len] = step[..., :kv_len]; [..., :kv_len-1] += previous.

    step_scores [L, B, W] (normalised), previous [L, B, kv_len-1] -> new cumulative [L, B, kv_len].
    """
    cum = step_scores[:, :, :kv_len].clone()
    cum[:, :, : kv_len - 1].add_(previous)
    return cum


def recent_count(budget: int, recent_ratio: float, kv_len: int) -> int:
    """h2o.py:498-499 / 541-542."""
    return min(max(1, int(budget * float(recent_ratio))), budget, kv_len)


def select_indices_batch(scores: torch.Tensor, *, budget: int, recent_ratio: float) -> torch.Tensor:
    """``select_h2o_indices_batch`` (h2o.py:519-563): scores [rows, kv_len] -> int64 [rows, budget].

    Heavy hitters = first (budget - recent) of a STABLE DESCENDING argsort over
    [0, kv_len - recent); recent suffix always kept; result ascending.
    """
    rows, kv_len = scores.shape
    if kv_len <= budget:
        return torch.arange(kv_len, dtype=torch.long).expand(rows, -1)
    rc = recent_count(budget, recent_ratio, kv_len)
    heavy = budget - rc
    start = kv_len - rc
    recent = torch.arange(start, kv_len, dtype=torch.long).expand(rows, -1)
    if heavy == 0:
        return recent
    top = torch.argsort(scores[:, :start], dim=1, descending=True, stable=True)[:, : min(heavy, start)]
    keep = torch.cat((top, recent), dim=1)
    assert keep.shape[1] == budget
    return torch.sort(keep, dim=1).values


def select_indices(scores: torch.Tensor, *, budget: int, recent_ratio: float) -> torch.Tensor:
    """``select_h2o_indices`` (h2o.py:479-517), 1-D."""
    return select_indices_batch(scores[None, :], budget=budget, recent_ratio=recent_ratio)[0]


def compact_rows(slot_rows: torch.Tensor, scores: torch.Tensor, keep: torch.Tensor):
    """Index-only compaction of ``free_part_slots_batch_layers`` (snapkv.py:1743-1803) + the
    kept-score gather of h2o.py:1588: slot_rows/scores [rows, kv_len], keep [rows, budget] ascending.

    Returns (new_slot_rows [rows, kv_len] with a zeroed tail, kept_scores [rows, budget],
    dropped_slots [rows, kv_len-budget] in ascending logical order — the order they are pushed
    on the free stack, row after row).
    """
    rows, kv_len = slot_rows.shape
    budget = keep.shape[1]
    new_slots = slot_rows.gather(1, keep)
    mask = torch.ones_like(slot_rows, dtype=torch.bool)
    mask.scatter_(1, keep, False)
    dropped = slot_rows[mask].view(rows, kv_len - budget)
    out = torch.zeros_like(slot_rows)
    out[:, :budget] = new_slots
    return out, scores.gather(1, keep), dropped


def prefill_window_score(q, k_cache, slot_table, req_indices, start_loc, context_lens, prompt_cache_lens, score_starts,
                         score_ends, *, candidate_start: int = 0, num_recent_tokens: int = 0):
    """Token scores of a prefill chunk: ``prefill_score_fwd`` in probability mode
    (kernels/triton/prefill_score.py:70-428; wrapper 430-680) as H2O drives it (h2o.py:778-894).

    For row b the queries at physical positions [score_starts[b], score_ends[b]) attend causally to kv positions
    in [candidate_start, max(candidate_start, ctx - num_recent)); logits are q.k * D^-0.5, masked entries -1e20,
    every (query, head) row is softmax-normalised over the candidates (a row with no candidate contributes 0);
    score[b][kv] = max over q heads of sum over queries of the probabilities / max(score_end - score_start, 1).
    Returns fp32 [B, max(context_lens)] (zeros outside the candidate range)."""
    bsz = int(req_indices.numel())
    hq, d = int(q.shape[1]), int(q.shape[2])
    hkv = int(k_cache.shape[1])
    g = hq // hkv
    out = torch.zeros((bsz, int(context_lens.max())), dtype=torch.float32)
    scale = float(d) ** -0.5
    for b in range(bsz):
        ctx, cache_len = int(context_lens[b]), int(prompt_cache_lens[b])
        chunk_len = ctx - cache_len
        qs, qe = int(score_starts[b]), int(score_ends[b])
        ce = max(candidate_start, ctx - num_recent_tokens)
        if ce <= candidate_start:
            continue
        q_abs = torch.arange(qs, qe)
        q_rel = q_abs - cache_len
        ok = (q_rel >= 0) & (q_rel < chunk_len)
        q_abs, q_rel = q_abs[ok], q_rel[ok]
        if q_abs.numel() == 0:
            continue
        kv_pos = torch.arange(candidate_start, ce)
        slots = slot_table[int(req_indices[b]), candidate_start:ce].long()
        k = k_cache[slots].float()                                  # [n, Hkv, D]
        qv = q[int(start_loc[b]) + q_rel].float()                   # [nq, Hq, D]
        kk = k.repeat_interleave(g, dim=1)                          # [n, Hq, D]
        logits = torch.einsum("qhd,nhd->hqn", qv, kk) * scale
        causal = q_abs[:, None] >= kv_pos[None, :]
        logits = torch.where(causal[None], logits, torch.full_like(logits, -1.0e20))
        m = logits.max(dim=-1, keepdim=True).values
        p = torch.where(causal[None], torch.exp(logits - m), torch.zeros_like(logits))
        l = p.sum(dim=-1, keepdim=True)
        p = p / torch.where(l > 0, l, torch.ones_like(l))
        head_score = p.sum(dim=1) / float(max(qe - qs, 1))          # [Hq, n]
        out[b, candidate_start:ce] = head_score.max(dim=0).values
    return out


def accumulate_prefill(previous, step_score, new_len: int, weight: float):
    """``H2OCacheManager._accumulate_score`` (h2o.py:611-626): zero-extend the previous cumulative vector to
    ``new_len`` and add ``weight * step_score[:new_len]``."""
    cum = torch.zeros(new_len, dtype=torch.float32)
    if previous is not None:
        cum[: previous.numel()] = previous.float()
    return cum + float(weight) * step_score[:new_len].float()
