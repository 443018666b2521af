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
