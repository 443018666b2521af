"""Torch-tensor front end of the C ABI (include/svb200.h).

Each function validates devices / dtypes / contiguity the way the reference's Triton wrappers
do (fail fast, kernels/triton/gqa_flash_decoding_stage1.py:298-352), passes raw device pointers
to libsvb200.so on the CURRENT torch stream and converts non-zero return codes into
RuntimeError.  No function here computes anything on the host and none has a fallback.
"""
from __future__ import annotations

import ctypes
import math

import torch

from . import _native
from ._native import check, current_stream_ptr, ptr, require_cuda

HEAD_DIM = 128
PAGE = 16


def _i32(t: torch.Tensor, name: str) -> None:
    require_cuda(t, name, torch.int32)
    if t.dim() >= 1 and t.stride(-1) != 1:
        raise ValueError(f"{name} must be contiguous in its last dimension, got strides {t.stride()}.")


def _rows_stride(t: torch.Tensor, name: str) -> int:
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be 2-D with a contiguous last dimension, got {tuple(t.shape)}/{t.stride()}.")
    return int(t.stride(0))


def _kv_layer(t: torch.Tensor, name: str) -> tuple[int, int]:
    """[slots, Hkv, 128] bf16 contiguous -> (num_slots, Hkv)."""
    require_cuda(t, name, torch.bfloat16)
    if t.dim() != 3 or t.shape[2] != HEAD_DIM or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous [slots, Hkv, {HEAD_DIM}] tensor, got {tuple(t.shape)}/{t.stride()}.")
    return int(t.shape[0]), int(t.shape[1])


def _q(t: torch.Tensor, name: str) -> tuple[int, int, int]:
    """[B, Hq, 128] bf16 with contiguous (head, dim) -> (B, Hq, batch stride)."""
    require_cuda(t, name, torch.bfloat16)
    if t.dim() != 3 or t.shape[2] != HEAD_DIM or t.stride(2) != 1 or t.stride(1) != HEAD_DIM:
        raise ValueError(f"{name} must be [B, H, {HEAD_DIM}] with contiguous heads, got {tuple(t.shape)}/{t.stride()}.")
    return int(t.shape[0]), int(t.shape[1]), int(t.stride(0))


def bounds_head_split(hkv: int) -> int:
    """Ring stages per 16-page unit: the smallest divisor of Hkv leaving <= 4 heads (csrc/bounds_layout.cuh)."""
    hkv = int(hkv)
    for d in range(1, hkv):
        if hkv % d == 0 and hkv // d <= 4:
            return d
    return 1 if hkv <= 4 else hkv


def bounds_shape(num_chunks: int, hkv: int) -> tuple[int, ...]:
    """One layer of Quest bounds: ``[chunks, hs, 16 pages, 2 (max, min), Hkv/hs, 128]`` (bounds_layout.cuh)."""
    hs = bounds_head_split(hkv)
    return (int(num_chunks), hs, PAGE, 2, int(hkv) // hs, HEAD_DIM)


def alloc_bounds(num_chunks: int, hkv: int, device, layers: int | None = None) -> torch.Tensor:
    shape = bounds_shape(num_chunks, hkv)
    return torch.zeros(shape if layers is None else (int(layers), *shape), dtype=torch.bfloat16, device=device)


def _bounds_layer(t: torch.Tensor, hkv: int, name: str = "bounds_layer") -> int:
    """Validates one layer of bounds chunks -> num_chunks."""
    require_cuda(t, name, torch.bfloat16)
    if t.dim() != 6 or tuple(t.shape[1:]) != bounds_shape(0, hkv)[1:] or not t.is_contiguous():
        raise ValueError(f"{name} must be a contiguous [chunks, {', '.join(map(str, bounds_shape(0, hkv)[1:]))}] tensor "
                         f"(ops.alloc_bounds), got {tuple(t.shape)}.")
    return int(t.shape[0])


def _swizzle_index(page_in_chunk: torch.Tensor) -> torch.Tensor:
    """[P] -> [P, 16]: position of 16-byte segment s of a row of page p is s ^ (p & 7)."""
    seg = torch.arange(16, device=page_in_chunk.device)
    return seg[None, :] ^ (page_in_chunk[:, None] & 7)


def pack_bounds(page_major: torch.Tensor, page_rec: torch.Tensor, bounds_layer: torch.Tensor) -> torch.Tensor:
    """Import page-major bounds ``[P, 2, Hkv, 128]`` (the reference's ``metadata_cache`` order, indexed by
    physical page) into this provider's chunk layout; pages with ``page_rec < 0`` are skipped.  Host-side
    indexing utility (tests, migration of an existing cache), not a kernel of the hot path."""
    p_total, _, hkv, _ = page_major.shape
    hs = bounds_head_split(hkv)
    hh = hkv // hs
    rec = page_rec.to(torch.long)
    pages = torch.nonzero(rec >= 0, as_tuple=False).flatten()
    rec = rec[pages]
    chunk, pg = rec >> 4, rec & 15
    x = page_major[pages].reshape(len(pages), 2, hs, hh, 16, 8)
    idx = _swizzle_index(pg)[:, None, None, None, :, None].expand(len(pages), 2, hs, hh, 16, 8)
    x = x.gather(4, idx)  # out[.., s, :] = in[.., s ^ (p & 7), :]
    bounds_layer[chunk, :, pg] = x.permute(0, 2, 1, 3, 4, 5).reshape(len(pages), hs, 2, hh, HEAD_DIM)
    return bounds_layer


def unpack_bounds(bounds_layer: torch.Tensor, page_rec: torch.Tensor) -> torch.Tensor:
    """Inverse of ``pack_bounds``: ``[P, 2, Hkv, 128]`` indexed by physical page (zeros where unassigned)."""
    _, hs, _, _, hh, _ = bounds_layer.shape
    hkv = hs * hh
    out = torch.zeros((int(page_rec.numel()), 2, hkv, HEAD_DIM), dtype=bounds_layer.dtype, device=bounds_layer.device)
    rec = page_rec.to(torch.long)
    pages = torch.nonzero(rec >= 0, as_tuple=False).flatten()
    rec = rec[pages]
    chunk, pg = rec >> 4, rec & 15
    x = bounds_layer[chunk, :, pg].reshape(len(pages), hs, 2, hh, 16, 8)
    idx = _swizzle_index(pg)[:, None, None, None, :, None].expand(len(pages), hs, 2, hh, 16, 8)
    x = x.gather(4, idx)
    out[pages] = x.permute(0, 2, 1, 3, 4, 5).reshape(len(pages), 2, hkv, HEAD_DIM)
    return out


def sm_count() -> int:
    return _native.sm_count()


# ----------------------------------------------------------------------------- Quest bounds
def quest_page_minmax(k_cache: torch.Tensor, bounds_layer: torch.Tensor, page_rec: torch.Tensor,
                      slot_mapping: torch.Tensor) -> None:
    """Rebuild the bounds of every page completed by ``slot_mapping`` (svb200_quest_page_minmax)."""
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    num_chunks = _bounds_layer(bounds_layer, hkv)
    _i32(slot_mapping, "slot_mapping"); _page_rec(page_rec, num_slots)
    lib = _native.load()
    check(lib.svb200_quest_page_minmax(ptr(k_cache), ptr(bounds_layer), num_chunks, ptr(page_rec), ptr(slot_mapping),
                                       int(slot_mapping.numel()), hkv, num_slots,
                                       current_stream_ptr(k_cache.device)), "svb200_quest_page_minmax")


def _page_rec(page_rec: torch.Tensor, num_slots: int) -> None:
    _i32(page_rec, "page_rec")
    if page_rec.dim() != 1 or int(page_rec.numel()) * PAGE < num_slots:
        raise ValueError(f"page_rec must be a 1-D int32 tensor with one entry per physical page "
                         f"(>= {num_slots // PAGE}), got {tuple(page_rec.shape)}.")


def quest_decode_page_complete(k_cache_all: torch.Tensor, bounds_all: torch.Tensor, page_rec: torch.Tensor,
                               page_table: torch.Tensor, req_rows: torch.Tensor, seq_lens: torch.Tensor) -> None:
    """All-layer refresh of pages completed by this decode step (svb200_quest_decode_page_complete).

    k_cache_all [L, slots, Hkv, 128], bounds_all [L, chunks, ...] (ops.alloc_bounds(layers=L))."""
    require_cuda(k_cache_all, "k_cache_all", torch.bfloat16)
    require_cuda(bounds_all, "bounds_all", torch.bfloat16)
    if k_cache_all.dim() != 4 or not k_cache_all.is_contiguous() or k_cache_all.shape[3] != HEAD_DIM:
        raise ValueError(f"k_cache_all must be contiguous [L, slots, Hkv, {HEAD_DIM}], got {tuple(k_cache_all.shape)}.")
    layers, num_slots, hkv, _ = (int(x) for x in k_cache_all.shape)
    if bounds_all.dim() != 7 or int(bounds_all.shape[0]) != layers:
        raise ValueError(f"bounds_all must be [L={layers}, chunks, ...] (ops.alloc_bounds), got {tuple(bounds_all.shape)}.")
    num_chunks = _bounds_layer(bounds_all[0], hkv, "bounds_all[0]")
    if not bounds_all.is_contiguous():
        raise ValueError("bounds_all must be contiguous.")
    _i32(page_table, "page_table"); _i32(req_rows, "req_rows"); _i32(seq_lens, "seq_lens"); _page_rec(page_rec, num_slots)
    lib = _native.load()
    check(lib.svb200_quest_decode_page_complete(
        ptr(k_cache_all), ptr(bounds_all), num_chunks, ptr(page_rec), ptr(page_table),
        _rows_stride(page_table, "page_table"), ptr(req_rows), ptr(seq_lens), int(req_rows.numel()), layers, hkv,
        num_slots, num_slots // PAGE, current_stream_ptr(k_cache_all.device)), "svb200_quest_decode_page_complete")


# ----------------------------------------------------------------------------- Quest select
def score_row_stride(n_cols: int) -> int:
    return (int(n_cols) + 7) // 8 * 8


def quest_page_score(q: torch.Tensor, bounds_layer: torch.Tensor, unit_table: torch.Tensor,
                     req_rows: torch.Tensor, context_lens: torch.Tensor, n_cols: int,
                     num_kv_heads: int, scores: torch.Tensor | None = None,
                     sched: torch.Tensor | None = None) -> torch.Tensor:
    """bf16 page scores [B, stride>=n_cols] (svb200_quest_page_score).  ``sched``: sched_scratch(B)."""
    bsz, hq, qs = _q(q, "q")
    if sched is None:
        sched = sched_scratch(bsz, q.device)
    _i32(sched, "sched")
    if sched.numel() < sched_scratch_ints(bsz):
        raise ValueError(f"sched needs >= {sched_scratch_ints(bsz)} int32 entries, got {sched.numel()}.")
    num_chunks = _bounds_layer(bounds_layer, num_kv_heads)
    _i32(unit_table, "unit_table"); _i32(req_rows, "req_rows"); _i32(context_lens, "context_lens")
    if scores is None:
        scores = torch.empty((bsz, score_row_stride(n_cols)), dtype=torch.bfloat16, device=q.device)
    require_cuda(scores, "scores", torch.bfloat16)
    lib = _native.load()
    check(lib.svb200_quest_page_score(
        ptr(q), qs, ptr(bounds_layer), num_chunks, ptr(unit_table), _rows_stride(unit_table, "unit_table"),
        ptr(req_rows), ptr(context_lens), ptr(scores), _rows_stride(scores, "scores"), bsz, hq,
        int(num_kv_heads), int(n_cols), ptr(sched), current_stream_ptr(q.device)), "svb200_quest_page_score")
    return scores


def quest_select_pack(scores: torch.Tensor, n_cols: int, page_table: torch.Tensor, req_rows: torch.Tensor,
                      context_lens: torch.Tensor, prev_budget: int, packed_slots: torch.Tensor,
                      sparse_lens: torch.Tensor, selected_pages: torch.Tensor | None = None,
                      dense_token_table: torch.Tensor | None = None, token_budget: int = 0,
                      page_budget_base: int = 0, out_width: int = 0) -> None:
    """Top-k + last page + slot packing (svb200_quest_select_pack)."""
    require_cuda(scores, "scores", torch.bfloat16)
    _i32(page_table, "page_table"); _i32(req_rows, "req_rows"); _i32(context_lens, "context_lens")
    _i32(packed_slots, "packed_slots"); _i32(sparse_lens, "sparse_lens")
    bsz = int(req_rows.numel())
    if selected_pages is not None:
        _i32(selected_pages, "selected_pages")
        if tuple(selected_pages.shape) != (bsz, prev_budget + 1) or not selected_pages.is_contiguous():
            raise ValueError("selected_pages must be a contiguous [B, prev_budget+1] int32 tensor.")
    dense_stride = 0
    if dense_token_table is not None:
        _i32(dense_token_table, "dense_token_table")
        dense_stride = _rows_stride(dense_token_table, "dense_token_table")
    lib = _native.load()
    check(lib.svb200_quest_select_pack(
        ptr(scores), _rows_stride(scores, "scores"), int(n_cols), ptr(page_table),
        _rows_stride(page_table, "page_table"), ptr(req_rows), ptr(context_lens), bsz, int(prev_budget),
        ptr(packed_slots), _rows_stride(packed_slots, "packed_slots"), ptr(sparse_lens), ptr(selected_pages),
        ptr(dense_token_table), dense_stride, int(token_budget), int(page_budget_base), int(out_width),
        current_stream_ptr(scores.device)), "svb200_quest_select_pack")


class DecodeStore:
    """New K/V rows of a decode step (token b = batch row b) for the scoring launch to store itself."""

    __slots__ = ("key", "value", "k_cache", "v_cache", "slot_mapping", "refresh_bounds", "page_rec")

    def __init__(self, key, value, k_cache, v_cache, slot_mapping, refresh_bounds: bool, page_rec=None):
        self.key, self.value, self.k_cache, self.v_cache = key, value, k_cache, v_cache
        self.slot_mapping, self.refresh_bounds, self.page_rec = slot_mapping, bool(refresh_bounds), page_rec


STATIC_MAX_BATCH = 1024  # SVB200_STATIC_MAX_BATCH (include/svb200.h)


def quest_score_select_pack_static(q: torch.Tensor, bounds_layer: torch.Tensor, unit_table: torch.Tensor,
                                   page_table: torch.Tensor, req_rows: torch.Tensor, context_lens: torch.Tensor,
                                   n_cols: int, num_kv_heads: int, prev_budget: int, scores: torch.Tensor,
                                   packed_slots: torch.Tensor, sparse_lens: torch.Tensor,
                                   selected_pages: torch.Tensor | None = None) -> None:
    """svb200_quest_score_select_pack_static: `quest_score_select_pack` for callers that cannot allocate: the
    scheduler scratch is a static device array inside the library (batch <= STATIC_MAX_BATCH; calls that overlap
    on the GPU race on it).  No counterpart in the reference.  This wrapper allocates nothing either."""
    bsz, hq, qs = _q(q, "q")
    num_chunks = _bounds_layer(bounds_layer, num_kv_heads)
    require_cuda(scores, "scores", torch.bfloat16)
    for t, n in ((unit_table, "unit_table"), (page_table, "page_table"), (req_rows, "req_rows"),
                 (context_lens, "context_lens"), (packed_slots, "packed_slots"), (sparse_lens, "sparse_lens")):
        _i32(t, n)
    if selected_pages is not None:
        _i32(selected_pages, "selected_pages")
    lib = _native.load()
    check(lib.svb200_quest_score_select_pack_static(
        ptr(q), qs, ptr(bounds_layer), num_chunks, ptr(unit_table), _rows_stride(unit_table, "unit_table"),
        ptr(page_table), _rows_stride(page_table, "page_table"), ptr(req_rows), ptr(context_lens), ptr(scores),
        _rows_stride(scores, "scores"), bsz, hq, int(num_kv_heads), int(n_cols), int(prev_budget), ptr(packed_slots),
        _rows_stride(packed_slots, "packed_slots"), ptr(sparse_lens), ptr(selected_pages),
        current_stream_ptr(q.device)), "svb200_quest_score_select_pack_static")


def quest_score_select_pack(q: torch.Tensor, bounds_layer: torch.Tensor, unit_table: torch.Tensor,
                            page_table: torch.Tensor, req_rows: torch.Tensor, context_lens: torch.Tensor,
                            n_cols: int, num_kv_heads: int,
                            prev_budget: int, scores: torch.Tensor, packed_slots: torch.Tensor,
                            sparse_lens: torch.Tensor, tickets: torch.Tensor,
                            selected_pages: torch.Tensor | None = None, store: DecodeStore | None = None) -> None:
    """One-launch score + select + pack for long-text batches (svb200_quest_score_select_pack); with
    ``store`` the launch also writes the step's K/V rows (svb200_quest_store_score_select_pack)."""
    bsz, hq, qs = _q(q, "q")
    num_chunks = _bounds_layer(bounds_layer, num_kv_heads)
    require_cuda(scores, "scores", torch.bfloat16)
    for t, n in ((unit_table, "unit_table"), (page_table, "page_table"), (req_rows, "req_rows"),
                 (context_lens, "context_lens"), (packed_slots, "packed_slots"), (sparse_lens, "sparse_lens"),
                 (tickets, "tickets")):
        _i32(t, n)
    if tickets.numel() < sched_scratch_ints(bsz):
        raise ValueError(f"tickets needs >= {sched_scratch_ints(bsz)} int32 entries (sched_scratch), got {tickets.numel()}.")
    if selected_pages is not None:
        _i32(selected_pages, "selected_pages")
    lib = _native.load()
    score_args = (ptr(q), qs, ptr(bounds_layer), num_chunks, ptr(unit_table), _rows_stride(unit_table, "unit_table"),
                  ptr(page_table), _rows_stride(page_table, "page_table"),
                  ptr(req_rows), ptr(context_lens), ptr(scores), _rows_stride(scores, "scores"), bsz, hq,
                  int(num_kv_heads), int(n_cols), int(prev_budget), ptr(packed_slots),
                  _rows_stride(packed_slots, "packed_slots"), ptr(sparse_lens), ptr(selected_pages), ptr(tickets),
                  current_stream_ptr(q.device))
    if store is None:
        check(lib.svb200_quest_score_select_pack(*score_args), "svb200_quest_score_select_pack")
        return
    num_slots, hkv = _kv_layer(store.k_cache, "k_cache")
    _kv_layer(store.v_cache, "v_cache")
    if hkv != int(num_kv_heads):
        raise ValueError(f"store caches hold {hkv} KV heads, scoring uses {num_kv_heads}.")
    for t, n in ((store.key, "key"), (store.value, "value")):
        require_cuda(t, n, torch.bfloat16)
        if t.dim() != 3 or tuple(t.shape) != (bsz, hkv, HEAD_DIM) or t.stride(2) != 1 or t.stride(1) != HEAD_DIM:
            raise ValueError(f"{n} must be [{bsz}, {hkv}, {HEAD_DIM}] with contiguous heads, got {tuple(t.shape)}/{t.stride()}.")
    _i32(store.slot_mapping, "slot_mapping")
    if int(store.slot_mapping.numel()) != bsz:
        raise ValueError(f"slot_mapping must hold one slot per batch row ({bsz}), got {store.slot_mapping.numel()}.")
    if store.refresh_bounds:
        if store.page_rec is None:
            raise ValueError("a store that refreshes bounds needs page_rec.")
        _page_rec(store.page_rec, num_slots)
    check(lib.svb200_quest_store_score_select_pack(
        ptr(store.key), int(store.key.stride(0)), ptr(store.value), int(store.value.stride(0)), ptr(store.k_cache),
        ptr(store.v_cache), ptr(store.slot_mapping), num_slots, int(store.refresh_bounds), ptr(store.page_rec),
        *score_args), "svb200_quest_store_score_select_pack")


def sched_scratch_ints(rows: int) -> int:
    """int32 entries of the scheduler scratch of the persistent score kernels for ``rows`` batch rows."""
    return int(_native.load().svb200_sched_scratch_ints(int(rows)))


def sched_scratch(rows: int, device) -> torch.Tensor:
    """Zeroed scheduler scratch (allocate once per stream; launches leave it zeroed)."""
    return torch.zeros((sched_scratch_ints(rows),), dtype=torch.int32, device=device)


def set_tuning(which: int, value: int) -> int:
    """Experiment knob (svb200_set_tuning); returns the previous value."""
    return int(_native.load().svb200_set_tuning(int(which), int(value)))


def set_max_ctas(n: int) -> int:
    """Cap the persistent kernels' grids at ``n`` CTAs (<= 0: one per SM); returns the previous cap."""
    return int(_native.load().svb200_set_max_ctas(int(n)))


# ----------------------------------------------------------------------------- attention
def attn_workspace(batch: int, num_q_heads: int, num_kv_heads: int, device, max_ctas: int | None = None) -> torch.Tensor:
    """Zeroed scratch of svb200_workspace_bytes (arrival tickets + per-SM fp32 partials).

    Allocate it once per (batch, heads) shape; kernels leave the ticket part zeroed.  ``max_ctas`` sizes it
    for fewer CTAs than SMs (the launch then uses at most that many; tests use it to vary the decomposition)."""
    lib = _native.load()
    nbytes = int(lib.svb200_workspace_bytes(int(batch), int(num_q_heads), int(num_kv_heads)))
    if nbytes <= 0:
        raise ValueError(f"svb200_workspace_bytes({batch}, {num_q_heads}, {num_kv_heads}) = {nbytes}.")
    if max_ctas is not None:
        sms = int(lib.svb200_sm_count())
        tickets = (int(batch) * int(num_kv_heads) * 4 + 255) // 256 * 256
        per_cta = (nbytes - tickets) // sms
        nbytes = tickets + max(1, min(int(max_ctas), sms)) * per_cta
    return torch.zeros(((nbytes + 3) // 4,), dtype=torch.int32, device=device)


def sparse_decode_attn(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, slot_table: torch.Tensor,
                       req_rows: torch.Tensor, seq_lens: torch.Tensor, *, out: torch.Tensor,
                       workspace: torch.Tensor, max_len: int | None = None,
                       attn_score: torch.Tensor | None = None, heads_per_cta: int = 0,
                       sm_scale: float | None = None, comm=None, gathered: torch.Tensor | None = None) -> torch.Tensor:
    """Stream-K sparse paged GQA decode attention with fused combine (svb200_sparse_decode_attn).

    ``comm`` (a ``parallel.PeerComm``) + ``gathered`` (a ``[B, world*Hq, 128]`` view INSIDE ``comm``'s local data
    buffer): the epilogue also stores this rank's rows into every rank's gathered buffer and the launch completes
    when all ranks' rows have arrived (svb200_sparse_decode_attn_tp)."""
    bsz, hq, qs = _q(q, "q")
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    if _kv_layer(v_cache, "v_cache") != (num_slots, hkv):
        raise ValueError("k_cache and v_cache must have identical shapes.")
    ob, oh, os_ = _q(out, "out")
    if (ob, oh) != (bsz, hq):
        raise ValueError(f"out must be [B, Hq, {HEAD_DIM}] like q, got {tuple(out.shape)}.")
    _i32(slot_table, "slot_table"); _i32(req_rows, "req_rows"); _i32(seq_lens, "seq_lens")
    if int(req_rows.numel()) != bsz or int(seq_lens.numel()) != bsz:
        raise ValueError("req_rows / seq_lens must have one entry per q row.")
    table_stride = _rows_stride(slot_table, "slot_table")
    width = int(slot_table.shape[-1])
    max_len = width if max_len is None else int(max_len)
    if max_len < 0 or max_len > width:
        raise ValueError(f"max_len={max_len} must be in [0, slot table width {width}].")
    _i32(workspace, "workspace")
    score_stride = 0
    if attn_score is not None:
        require_cuda(attn_score, "attn_score", torch.float32)
        if attn_score.dim() != 2 or attn_score.shape[0] < bsz or attn_score.stride(1) != 1 \
                or attn_score.shape[1] < max_len:
            raise ValueError(f"attn_score must be fp32 [B, W >= max_len] with contiguous W, got {tuple(attn_score.shape)}.")
        score_stride = int(attn_score.stride(0))
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(HEAD_DIM)
    lib = _native.load()
    args = (ptr(q), qs, ptr(k_cache), ptr(v_cache), num_slots, ptr(slot_table), table_stride, max_len,
            ptr(req_rows), ptr(seq_lens), ptr(out), os_, ptr(attn_score), score_stride, bsz, hq, hkv,
            int(heads_per_cta), ptr(workspace), int(workspace.numel()) * 4, float(sm_scale))
    if comm is None or comm.world <= 1:
        check(lib.svb200_sparse_decode_attn(*args, current_stream_ptr(q.device)), "svb200_sparse_decode_attn")
        return out
    if gathered is None:
        raise ValueError("the tensor-parallel epilogue needs the gathered output view.")
    require_cuda(gathered, "gathered", torch.bfloat16)
    gb, gh, gs = _q(gathered, "gathered")
    if (gb, gh) != (bsz, comm.world * hq):
        raise ValueError(f"gathered must be [B={bsz}, world*Hq={comm.world * hq}, {HEAD_DIM}], got {tuple(gathered.shape)}.")
    check(lib.svb200_sparse_decode_attn_tp(*args, comm.struct_ptr(), comm.offset_of(gathered), gs,
                                           current_stream_ptr(q.device)), "svb200_sparse_decode_attn_tp")
    return out


def quest_decode_layer(q: torch.Tensor, bounds_layer: torch.Tensor, unit_table: torch.Tensor, page_table: torch.Tensor,
                       req_rows: torch.Tensor, context_lens: torch.Tensor, n_cols: int, num_kv_heads: int,
                       prev_budget: int, scores: torch.Tensor, packed_slots: torch.Tensor, sparse_lens: torch.Tensor,
                       sched_scratch: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, *, out: torch.Tensor,
                       workspace: torch.Tensor, selected_pages: torch.Tensor | None = None,
                       store: DecodeStore | None = None, sm_scale: float | None = None, comm=None,
                       gathered: torch.Tensor | None = None) -> torch.Tensor:
    """One launch per sparse decode layer (svb200_quest_decode_layer): (store +) score + select + pack, then the
    sparse decode attention over the packed slots inside the same launch.  Same results as
    `quest_score_select_pack` followed by `sparse_decode_attn`.  Needs n_cols > 0 and <= 8 q heads per KV head."""
    bsz, hq, qs = _q(q, "q")
    num_chunks = _bounds_layer(bounds_layer, num_kv_heads)
    require_cuda(scores, "scores", torch.bfloat16)
    for t, n in ((unit_table, "unit_table"), (page_table, "page_table"), (req_rows, "req_rows"),
                 (context_lens, "context_lens"), (packed_slots, "packed_slots"), (sparse_lens, "sparse_lens"),
                 (sched_scratch, "sched_scratch"), (workspace, "workspace")):
        _i32(t, n)
    if sched_scratch.numel() < sched_scratch_ints(bsz):
        raise ValueError(f"sched_scratch needs >= {sched_scratch_ints(bsz)} int32 entries, got {sched_scratch.numel()}.")
    if selected_pages is not None:
        _i32(selected_pages, "selected_pages")
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    if _kv_layer(v_cache, "v_cache") != (num_slots, hkv) or hkv != int(num_kv_heads):
        raise ValueError("k_cache / v_cache must have identical shapes with num_kv_heads heads.")
    ob, oh, os_ = _q(out, "out")
    if (ob, oh) != (bsz, hq):
        raise ValueError(f"out must be [B, Hq, {HEAD_DIM}] like q, got {tuple(out.shape)}.")
    key = value = slot_mapping = page_rec = None
    k_stride = v_stride = refresh = 0
    if store is not None:
        if store.k_cache.data_ptr() != k_cache.data_ptr() or store.v_cache.data_ptr() != v_cache.data_ptr():
            raise ValueError("the fused store must write the caches the attention reads.")
        for t, n in ((store.key, "key"), (store.value, "value")):
            require_cuda(t, n, torch.bfloat16)
            if t.dim() != 3 or tuple(t.shape) != (bsz, hkv, HEAD_DIM) or t.stride(2) != 1 or t.stride(1) != HEAD_DIM:
                raise ValueError(f"{n} must be [{bsz}, {hkv}, {HEAD_DIM}] with contiguous heads, got {tuple(t.shape)}/{t.stride()}.")
        _i32(store.slot_mapping, "slot_mapping")
        if int(store.slot_mapping.numel()) != bsz:
            raise ValueError(f"slot_mapping must hold one slot per batch row ({bsz}), got {store.slot_mapping.numel()}.")
        if store.refresh_bounds:
            if store.page_rec is None:
                raise ValueError("a store that refreshes bounds needs page_rec.")
            _page_rec(store.page_rec, num_slots)
        key, value, slot_mapping, page_rec = store.key, store.value, store.slot_mapping, store.page_rec
        k_stride, v_stride, refresh = int(store.key.stride(0)), int(store.value.stride(0)), int(store.refresh_bounds)
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(HEAD_DIM)
    push = comm is not None and comm.world > 1
    gather_off = gather_stride = 0
    if push:
        if gathered is None:
            raise ValueError("the tensor-parallel epilogue needs the gathered output view.")
        require_cuda(gathered, "gathered", torch.bfloat16)
        gb, gh, gather_stride = _q(gathered, "gathered")
        if (gb, gh) != (bsz, comm.world * hq):
            raise ValueError(f"gathered must be [B={bsz}, world*Hq={comm.world * hq}, {HEAD_DIM}], got {tuple(gathered.shape)}.")
        gather_off = comm.offset_of(gathered)
    lib = _native.load()
    check(lib.svb200_quest_decode_layer(
        ptr(key), k_stride, ptr(value), v_stride, ptr(slot_mapping), refresh, ptr(page_rec),
        ptr(k_cache), ptr(v_cache), num_slots,
        ptr(q), qs, ptr(bounds_layer), num_chunks, ptr(unit_table), _rows_stride(unit_table, "unit_table"),
        ptr(page_table), _rows_stride(page_table, "page_table"), ptr(req_rows), ptr(context_lens), ptr(scores),
        _rows_stride(scores, "scores"), bsz, hq, hkv, int(n_cols), int(prev_budget), ptr(packed_slots),
        _rows_stride(packed_slots, "packed_slots"), ptr(sparse_lens), ptr(selected_pages), ptr(sched_scratch),
        ptr(out), os_, ptr(workspace), int(workspace.numel()) * 4, float(sm_scale),
        comm.struct_ptr() if push else None, gather_off, gather_stride, current_stream_ptr(q.device)),
        "svb200_quest_decode_layer")
    return out


def allreduce_sum_bf16(comm, inp: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """One-shot NVLink SUM all-reduce (svb200_allreduce_sum_bf16): ``inp`` is a contiguous bf16 view INSIDE
    ``comm``'s local data buffer (every rank passes the view at the same offset), ``out`` a local bf16 tensor."""
    require_cuda(inp, "inp", torch.bfloat16)
    require_cuda(out, "out", torch.bfloat16)
    if not inp.is_contiguous() or not out.is_contiguous() or inp.numel() != out.numel() or inp.numel() % 8:
        raise ValueError("inp / out must be contiguous bf16 tensors of the same size, a multiple of 8 elements.")
    lib = _native.load()
    check(lib.svb200_allreduce_sum_bf16(comm.struct_ptr(), comm.offset_of(inp), int(inp.numel()), ptr(out),
                                        current_stream_ptr(inp.device)), "svb200_allreduce_sum_bf16")
    return out


# ----------------------------------------------------------------------------- H2O
def h2o_normalize_accumulate(score: torch.Tensor, scale: float, cumulative: torch.Tensor | None = None,
                             kv_lens: torch.Tensor | None = None, width: int | None = None,
                             cum_rows: torch.Tensor | None = None, refill: float = 0.0) -> torch.Tensor:
    """In-place ``softmax(score * scale)`` per row (+ fused cumulative update) (svb200_h2o_normalize_accumulate).
    ``cum_rows`` [B] int32: row of each batch entry inside ``cumulative`` (default: entry b -> row b);
    ``refill`` != 0: leave the score rows holding ``refill`` instead of the normalised values."""
    require_cuda(score, "score", torch.float32)
    if score.dim() != 2 or score.stride(1) != 1:
        raise ValueError(f"score must be fp32 [B, W] with contiguous W, got {tuple(score.shape)}/{score.stride()}.")
    bsz = int(score.shape[0])
    width = int(score.shape[1]) if width is None else int(width)
    cum_stride = 0
    if cumulative is not None:
        require_cuda(cumulative, "cumulative", torch.float32)
        if cumulative.dim() != 2 or cumulative.stride(1) != 1 or kv_lens is None or \
                (cum_rows is None and cumulative.shape[0] < bsz):
            raise ValueError("cumulative must be fp32 [rows, >=kv_len] and needs kv_lens.")
        _i32(kv_lens, "kv_lens")
        if cum_rows is not None:
            _i32(cum_rows, "cum_rows")
            if int(cum_rows.numel()) != bsz:
                raise ValueError("cum_rows must hold one row index per batch entry.")
        cum_stride = int(cumulative.stride(0))
    lib = _native.load()
    check(lib.svb200_h2o_normalize_accumulate(
        ptr(score), int(score.stride(0)), bsz, width, float(scale), ptr(cumulative), cum_stride,
        ptr(cum_rows) if cumulative is not None else None, ptr(kv_lens), float(refill),
        current_stream_ptr(score.device)), "svb200_h2o_normalize_accumulate")
    return score


def h2o_prepare_decode(meta: torch.Tensor, free_stack: torch.Tensor, slot_tables: torch.Tensor,
                       layer_slot_mapping: torch.Tensor, context_lens: torch.Tensor, req_indices: torch.Tensor,
                       slot_mapping: torch.Tensor, kv_lens: torch.Tensor, rows_out: torch.Tensor, batch: int) -> None:
    """One launch for the device side of an H2O decode step's preparation (svb200_h2o_prepare_decode)."""
    for t, n in ((meta, "meta"), (free_stack, "free_stack"), (slot_tables, "slot_tables"),
                 (layer_slot_mapping, "layer_slot_mapping"), (context_lens, "context_lens"), (req_indices, "req_indices"),
                 (slot_mapping, "slot_mapping"), (kv_lens, "kv_lens"), (rows_out, "rows_out")):
        _i32(t, n)
    if slot_tables.dim() != 3 or free_stack.dim() != 2 or layer_slot_mapping.dim() != 2 or meta.numel() < 2 * batch + 1:
        raise ValueError("h2o_prepare_decode: slot_tables [L, rows, W], free_stack [L, slots], layer_slot_mapping [L, rows], "
                         "meta [2B+1] expected.")
    layers = int(slot_tables.shape[0])
    lib = _native.load()
    check(lib.svb200_h2o_prepare_decode(
        ptr(meta), ptr(free_stack), int(free_stack.stride(0)), ptr(slot_tables), int(slot_tables.stride(0)),
        int(slot_tables.stride(1)), ptr(layer_slot_mapping), int(layer_slot_mapping.stride(0)), ptr(context_lens),
        ptr(req_indices), ptr(slot_mapping), ptr(kv_lens), ptr(rows_out), int(batch), layers,
        current_stream_ptr(slot_tables.device)), "svb200_h2o_prepare_decode")


def prefill_attn(q: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor, slot_table: torch.Tensor,
                 req_rows: torch.Tensor, start_loc: torch.Tensor, context_lens: torch.Tensor,
                 prompt_cache_lens: torch.Tensor, *, max_chunk_len: int, out: torch.Tensor | None = None,
                 sm_scale: float | None = None) -> torch.Tensor:
    """Dense causal prefill attention of a chunk over its paged row (svb200_prefill_attn)."""
    require_cuda(q, "q", torch.bfloat16)
    if q.dim() != 3 or q.shape[2] != HEAD_DIM or q.stride(2) != 1 or q.stride(1) != HEAD_DIM:
        raise ValueError(f"q must be [T, Hq, {HEAD_DIM}] with contiguous heads, got {tuple(q.shape)}/{q.stride()}.")
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    if _kv_layer(v_cache, "v_cache") != (num_slots, hkv):
        raise ValueError("k_cache and v_cache must have identical shapes.")
    for t, n in ((slot_table, "slot_table"), (req_rows, "req_rows"), (start_loc, "start_loc"),
                 (context_lens, "context_lens"), (prompt_cache_lens, "prompt_cache_lens")):
        _i32(t, n)
    if out is None:
        out = torch.empty_like(q)
    require_cuda(out, "out", torch.bfloat16)
    if out.shape != q.shape or out.stride(2) != 1 or out.stride(1) != HEAD_DIM:
        raise ValueError("out must have q's shape with contiguous heads.")
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(HEAD_DIM)
    lib = _native.load()
    check(lib.svb200_prefill_attn(
        ptr(q), int(q.stride(0)), ptr(k_cache), ptr(v_cache), num_slots, ptr(slot_table), _rows_stride(slot_table, "slot_table"),
        ptr(req_rows), ptr(start_loc), ptr(context_lens), ptr(prompt_cache_lens), ptr(out), int(out.stride(0)),
        int(req_rows.numel()), int(q.shape[1]), hkv, int(max_chunk_len), float(sm_scale),
        current_stream_ptr(q.device)), "svb200_prefill_attn")
    return out


def h2o_prefill_score(q: torch.Tensor, k_cache: torch.Tensor, slot_table: torch.Tensor, req_rows: torch.Tensor,
                      start_loc: torch.Tensor, context_lens: torch.Tensor, prompt_cache_lens: torch.Tensor,
                      score_starts: torch.Tensor, score_ends: torch.Tensor, score: torch.Tensor, *, max_query_len: int,
                      max_context_len: int, cumulative: torch.Tensor | None = None, cum_rows: torch.Tensor | None = None,
                      candidate_start: int = 0, num_recent_tokens: int = 0, sm_scale: float | None = None) -> torch.Tensor:
    """Prefill window token scores (+ fused cumulative update) (svb200_h2o_prefill_score)."""
    require_cuda(q, "q", torch.bfloat16)
    if q.dim() != 3 or q.shape[2] != HEAD_DIM or q.stride(2) != 1 or q.stride(1) != HEAD_DIM:
        raise ValueError(f"q must be [T, Hq, {HEAD_DIM}] with contiguous heads, got {tuple(q.shape)}/{q.stride()}.")
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    for t, n in ((slot_table, "slot_table"), (req_rows, "req_rows"), (start_loc, "start_loc"), (context_lens, "context_lens"),
                 (prompt_cache_lens, "prompt_cache_lens"), (score_starts, "score_starts"), (score_ends, "score_ends")):
        _i32(t, n)
    bsz = int(req_rows.numel())
    require_cuda(score, "score", torch.float32)
    if score.dim() != 2 or score.shape[0] < bsz or score.stride(1) != 1 or score.shape[1] < max_context_len:
        raise ValueError(f"score must be fp32 [B, >= max_context_len] with contiguous columns, got {tuple(score.shape)}.")
    cum_stride = 0
    if cumulative is not None:
        require_cuda(cumulative, "cumulative", torch.float32)
        if cumulative.dim() != 2 or cumulative.stride(1) != 1 or cum_rows is None:
            raise ValueError("cumulative must be fp32 [rows, >= max_context_len] and needs cum_rows.")
        _i32(cum_rows, "cum_rows")
        cum_stride = int(cumulative.stride(0))
    if sm_scale is None:
        sm_scale = 1.0 / math.sqrt(HEAD_DIM)
    lib = _native.load()
    check(lib.svb200_h2o_prefill_score(
        ptr(q), int(q.stride(0)), ptr(k_cache), num_slots, ptr(slot_table), _rows_stride(slot_table, "slot_table"),
        ptr(req_rows), ptr(start_loc), ptr(context_lens), ptr(prompt_cache_lens), ptr(score_starts), ptr(score_ends),
        ptr(score), int(score.stride(0)), ptr(cumulative), cum_stride, ptr(cum_rows) if cumulative is not None else None,
        bsz, int(q.shape[1]), hkv, int(max_query_len), int(max_context_len), int(candidate_start),
        int(num_recent_tokens), float(sm_scale), current_stream_ptr(q.device)), "svb200_h2o_prefill_score")
    return score


def h2o_select_compact(cumulative: torch.Tensor, cum_rows: torch.Tensor, slot_tables: torch.Tensor,
                       slot_rows: torch.Tensor, *, kv_len: int, budget: int, recent: int,
                       free_stack: torch.Tensor | None = None, free_rows: torch.Tensor | None = None,
                       keep_out: torch.Tensor | None = None, kv_lens: torch.Tensor | None = None) -> None:
    """Burst eviction: heavy hitters + recent suffix, index-only compaction (svb200_h2o_select_compact).

    ``cum_rows`` / ``slot_rows`` / ``free_rows`` are int64 ELEMENT offsets of each (layer,row) entry
    inside ``cumulative`` / ``slot_tables`` / ``free_stack``.  ``kv_lens`` [num_rows] int32: per-entry lengths
    (then ``kv_len`` is the longest); entries not longer than ``budget`` are left alone."""
    require_cuda(cumulative, "cumulative", torch.float32)
    _i32(slot_tables, "slot_tables")
    for t, n in ((cum_rows, "cum_rows"), (slot_rows, "slot_rows")):
        require_cuda(t, n, torch.int64)
    n_rows = int(cum_rows.numel())
    if int(slot_rows.numel()) != n_rows:
        raise ValueError("cum_rows and slot_rows must have the same number of entries.")
    if free_stack is not None:
        _i32(free_stack, "free_stack"); require_cuda(free_rows, "free_rows", torch.int64)
    if kv_lens is not None:
        _i32(kv_lens, "kv_lens")
        if int(kv_lens.numel()) != n_rows:
            raise ValueError("kv_lens must hold one length per (layer, row) entry.")
    if keep_out is not None:
        _i32(keep_out, "keep_out")
        if keep_out.dim() != 2 or int(keep_out.shape[0]) < n_rows or int(keep_out.shape[1]) != budget \
                or not keep_out.is_contiguous():
            raise ValueError("keep_out must be a contiguous [>= num_rows, budget] int32 tensor.")
    lib = _native.load()
    check(lib.svb200_h2o_select_compact(
        ptr(cumulative), ptr(cum_rows), ptr(slot_tables), ptr(slot_rows), ptr(free_stack), ptr(free_rows),
        ptr(keep_out), n_rows, ptr(kv_lens), int(kv_len), int(budget), int(recent),
        current_stream_ptr(cumulative.device)), "svb200_h2o_select_compact")


# ----------------------------------------------------------------------------- KV store
def store_kv_minmax(key: torch.Tensor, value: torch.Tensor, k_cache: torch.Tensor, v_cache: torch.Tensor,
                    slot_mapping: torch.Tensor, quest_bounds_layer: torch.Tensor | None = None,
                    page_rec: torch.Tensor | None = None) -> None:
    """Scatter new K/V rows (+ fused bounds refresh of completed pages) (svb200_store_kv_minmax)."""
    num_slots, hkv = _kv_layer(k_cache, "k_cache")
    _kv_layer(v_cache, "v_cache")
    for t, n in ((key, "key"), (value, "value")):
        require_cuda(t, n, torch.bfloat16)
        if t.dim() != 3 or tuple(t.shape[1:]) != (hkv, HEAD_DIM) or t.stride(2) != 1 or t.stride(1) != HEAD_DIM:
            raise ValueError(f"{n} must be [T, {hkv}, {HEAD_DIM}] with contiguous heads, got {tuple(t.shape)}/{t.stride()}.")
    _i32(slot_mapping, "slot_mapping")
    num_chunks = 0
    if quest_bounds_layer is not None:
        num_chunks = _bounds_layer(quest_bounds_layer, hkv, "quest_bounds_layer")
        if page_rec is None:
            raise ValueError("quest_bounds_layer needs page_rec (physical page -> bounds record).")
        _page_rec(page_rec, num_slots)
    lib = _native.load()
    check(lib.svb200_store_kv_minmax(
        ptr(key), int(key.stride(0)), ptr(value), int(value.stride(0)), ptr(k_cache), ptr(v_cache),
        ptr(slot_mapping), int(slot_mapping.numel()), hkv, num_slots, ptr(quest_bounds_layer),
        ptr(page_rec) if quest_bounds_layer is not None else None, num_chunks,
        current_stream_ptr(key.device)), "svb200_store_kv_minmax")
