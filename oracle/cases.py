"""Seeded synthetic inputs shared by oracle/gen_golden.py and tests/ (test infrastructure).

Inputs are regenerated from the seed with numpy's PCG64 *integer* stream only (bit-stable
across platforms), turned into approximately normal bf16 values with exact integer / IEEE
fp32 arithmetic (Irwin-Hall sum of 4 bytes), so the GPU box rebuilds exactly the tensors the
golden outputs were generated from; every golden file also stores a sha256 of the regenerated
tensors so a drift fails loudly instead of comparing against the wrong vectors.
Large tensors (K/V caches) therefore never need to be committed.
"""
from __future__ import annotations

import hashlib

import numpy as np
import torch

PAGE = 16

QUEST_CASES = {
    # long-text, ragged last pages incl. an exactly completed page (ctx % 16 == 0)
    "quest_small_long": dict(seed=11, bsz=3, hq=8, hkv=4, ctx_lens=[512, 500, 389],
                             max_context_len=512, token_budget=128, is_long_text=True),
    # Qwen2.5-7B head geometry (28q/4kv, group 7), context bucket larger than one row
    "quest_g7_bucket": dict(seed=12, bsz=2, hq=28, hkv=4, ctx_lens=[769, 1024],
                            max_context_len=1024, token_budget=256, is_long_text=True),
    # Llama geometry (32q/8kv), mixed short/long rows (dense_mask merge, not long-text)
    "quest_mixed_dense": dict(seed=13, bsz=3, hq=32, hkv=8, ctx_lens=[200, 600, 260],
                              max_context_len=1024, token_budget=256, is_long_text=False),
    # group 8 with one KV head (Qwen3-32B at TP=8)
    "quest_tp8_shard": dict(seed=14, bsz=2, hq=8, hkv=1, ctx_lens=[2000, 1999],
                            max_context_len=2048, token_budget=1024, is_long_text=True),
}

# BASELINE.json shapes at full size (one layer): inputs are regenerated from the seed with byte-wide draws
# (`compact`), only digests / scores / packed slots / attention outputs are committed.
QUEST_BASELINE_CASES = {
    # cfg1: tiny random-init Llama-style model (configs/debug/qwen3_tiny_random.json: 8 q / 4 kv heads), bs 1,
    # 4K context, budget 512 (sink 64 + decode 384 + recent 64), quest_skip_layers 0
    "quest_cfg1": dict(seed=41, bsz=1, hq=8, hkv=4, ctx_lens=[4091], max_context_len=4096, token_budget=512,
                       is_long_text=True, compact=True),
    # cfg2: Llama-3.1-8B shape (32 q / 8 kv), bs 1, 32K context, budget 2048
    "quest_cfg2": dict(seed=42, bsz=1, hq=32, hkv=8, ctx_lens=[32765], max_context_len=32768, token_budget=2048,
                       is_long_text=True, compact=True),
}

ATTN_CASES = {
    "attn_g7_ragged": dict(seed=21, bsz=3, hq=28, hkv=4, lens=[700, 513, 1], width=768,
                           num_slots=2048, with_score=False),
    "attn_g4_score": dict(seed=22, bsz=2, hq=32, hkv=8, lens=[320, 257], width=320,
                          num_slots=1024, with_score=True),
    "attn_g8_empty_row": dict(seed=23, bsz=3, hq=8, hkv=1, lens=[64, 0, 300], width=512,
                              num_slots=1024, with_score=True),
}

H2O_CASES = {
    "h2o_select_random": dict(seed=31, rows=6, kv_len=272, budget=256, ratio=0.5, ties=False),
    "h2o_select_ties": dict(seed=32, rows=5, kv_len=200, budget=96, ratio=0.25, ties=True),
}


# H2O prefill window scoring: rows = (physical prefix length before the chunk, chunk length); `window` = queries
# scored at the end of every chunk (h2o_prefill_score_window)
H2O_PREFILL_CASES = {
    "h2o_prefill_g4_w32": dict(seed=51, hq=8, hkv=2, rows=[(0, 200), (150, 100), (64, 20)], window=32, num_slots=1024),
    "h2o_prefill_g7_w32": dict(seed=52, hq=28, hkv=4, rows=[(96, 129), (0, 33)], window=32, num_slots=512),
    "h2o_prefill_g8_w128": dict(seed=53, hq=16, hkv=2, rows=[(300, 256), (0, 128)], window=128, num_slots=1024),
}


def _rng(seed: int) -> np.random.Generator:
    return np.random.Generator(np.random.PCG64(int(seed)))


def randn_bf16(rng: np.random.Generator, *shape: int, scale: float = 1.0, compact: bool = False) -> torch.Tensor:
    """~N(0, scale^2) bf16 tensor from integer draws: (sum of 4 bytes - 510) / 147.8 (exact in fp32).
    ``compact`` draws bytes (a different, equally bit-stable stream; 8x less memory for the large cases)."""
    u = rng.integers(0, 256, size=(*shape, 4), dtype=np.uint8 if compact else np.int64).sum(axis=-1, dtype=np.int64)
    x = (u - 510).astype(np.float32) * np.float32(scale) / np.float32(147.8)
    return torch.from_numpy(x).to(torch.bfloat16)


def randperm(rng: np.random.Generator, n: int) -> torch.Tensor:
    """Fisher-Yates from integer draws (does not rely on numpy's permutation algorithm)."""
    a = np.arange(n, dtype=np.int64)
    for i in range(n - 1, 0, -1):
        j = int(rng.integers(0, i + 1))
        a[i], a[j] = a[j], a[i]
    return torch.from_numpy(a)


def bf16_bits(t: torch.Tensor) -> np.ndarray:
    """bf16 tensor -> uint16 bit-pattern array."""
    assert t.dtype == torch.bfloat16
    return t.contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def bf16_from_bits(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(a.astype(np.uint16).view(np.int16).copy()).view(torch.bfloat16)


def digest(*tensors: torch.Tensor) -> str:
    h = hashlib.sha256()
    for t in tensors:
        t = t.contiguous()
        if t.dtype == torch.bfloat16:
            t = t.view(torch.int16)
        h.update(t.numpy().tobytes())
    return h.hexdigest()


def make_quest_inputs(*, seed, bsz, hq, hkv, ctx_lens, max_context_len, d=128, compact=False, **_):
    """Page tables with a shuffled physical page assignment, per-row K/V, q.

    Row 0 of the tables is left unused so that req_indices != arange is exercised.
    """
    gen = _rng(seed)
    rows = bsz + 1
    max_pages_per_row = (max_context_len + PAGE - 1) // PAGE
    used = sum((n + PAGE - 1) // PAGE for n in ctx_lens)
    num_pages = used + 3
    perm = randperm(gen, num_pages).to(torch.int32)
    page_table = torch.full((rows, max_pages_per_row), -1, dtype=torch.int32)
    token_table = torch.zeros((rows, max_context_len), dtype=torch.int32)
    req_indices = torch.arange(1, bsz + 1, dtype=torch.int32)
    ks, vs, slot_rows = [], [], []
    cursor = 0
    for b in range(bsz):
        n = int(ctx_lens[b])
        npg = (n + PAGE - 1) // PAGE
        pg = perm[cursor:cursor + npg]
        cursor += npg
        row = int(req_indices[b])
        page_table[row, :npg] = pg
        slots = (pg[:, None] * PAGE + torch.arange(PAGE, dtype=torch.int32)[None, :]).reshape(-1)[:n]
        token_table[row, :n] = slots
        ks.append(randn_bf16(gen, n, hkv, d, compact=compact))
        vs.append(randn_bf16(gen, n, hkv, d, compact=compact))
        slot_rows.append(slots)
    q = randn_bf16(gen, bsz, hq, d, compact=compact)
    return dict(num_pages=num_pages, max_pages_per_row=max_pages_per_row, page_table=page_table,
                token_table=token_table, req_indices=req_indices, k_rows=ks, v_rows=vs,
                slot_rows=slot_rows, q=q, context_lens=torch.tensor(ctx_lens, dtype=torch.int32))


def quest_fill_cache(inp, hkv, d=128):
    """K/V caches [num_pages*16, Hkv, D] with every row's tokens written at its slots."""
    k_cache = torch.zeros(inp["num_pages"] * PAGE, hkv, d, dtype=torch.bfloat16)
    v_cache = torch.zeros_like(k_cache)
    for k, v, slots in zip(inp["k_rows"], inp["v_rows"], inp["slot_rows"]):
        k_cache.index_copy_(0, slots.long(), k)
        v_cache.index_copy_(0, slots.long(), v)
    return k_cache, v_cache


def chunk_maps(page_table: torch.Tensor, num_pages: int, seed: int = 0, spare: int = 3):
    """Bounds-chunk assignment for a page table whose rows hold a prefix of valid pages (-1 after it):
    every (row, logical page // 16) gets a chunk id from a SHUFFLED pool (so the unit-table indirection is
    exercised), ``page_rec[physical page] = chunk * 16 + logical page % 16`` (csrc/bounds_layout.cuh).
    Returns (unit_table [rows, units] int32, page_rec [num_pages] int32, num_chunks)."""
    pt = page_table.cpu().numpy()
    rows, max_pages = pt.shape
    max_units = (max_pages + PAGE - 1) // PAGE
    num_chunks = rows * max_units + spare
    perm = np.random.Generator(np.random.PCG64(seed)).permutation(num_chunks).astype(np.int32)
    unit_table = np.full((rows, max_units), -1, dtype=np.int32)
    page_rec = np.full((num_pages,), -1, dtype=np.int32)
    k = 0
    for r in range(rows):
        npg = int((pt[r] >= 0).sum())
        nu = (npg + PAGE - 1) // PAGE
        unit_table[r, :nu] = perm[k: k + nu]
        k += nu
        j = np.arange(npg)
        page_rec[pt[r, :npg]] = unit_table[r, j // PAGE] * PAGE + j % PAGE
    return torch.from_numpy(unit_table), torch.from_numpy(page_rec), num_chunks


def make_attn_inputs(*, seed, bsz, hq, hkv, lens, width, num_slots, d=128, **_):
    gen = _rng(seed)
    q = randn_bf16(gen, bsz, hq, d)
    k_cache = randn_bf16(gen, num_slots, hkv, d)
    v_cache = randn_bf16(gen, num_slots, hkv, d)
    rows = bsz + 1
    table = torch.zeros((rows, width), dtype=torch.int32)
    for r in range(rows):
        table[r] = randperm(gen, num_slots)[:width].to(torch.int32)
    req = torch.arange(bsz, 0, -1, dtype=torch.int32)  # reversed rows
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, slot_table=table, req_indices=req,
                seq_lens=torch.tensor(lens, dtype=torch.int32))


def make_h2o_scores(*, seed, rows, kv_len, ties, **_):
    gen = _rng(seed)
    scores = torch.from_numpy(gen.integers(0, 1 << 24, size=(rows, kv_len)).astype(np.float32) / np.float32(1 << 24))
    if ties:
        scores = (scores * 8).floor() / 8  # massive ties: the stable order must decide
    return scores


def make_h2o_prefill_inputs(*, seed, hq, hkv, rows, window, num_slots, d=128, **_):
    """A prefill chunk per row on top of a compressed physical prefix: q of the chunk tokens (all rows
    concatenated), K cache, shuffled slot tables, and the score ranges of prefill_score_ranges (h2o.py:749-776)."""
    gen = _rng(seed)
    bsz = len(rows)
    k_cache = randn_bf16(gen, num_slots, hkv, d)
    ctx = [c + n for c, n in rows]
    width = max(ctx)
    table = torch.zeros((bsz + 1, width), dtype=torch.int32)
    perm = randperm(gen, num_slots).to(torch.int32)
    cursor = 0
    req = torch.arange(bsz, 0, -1, dtype=torch.int32)  # reversed table rows
    for b in range(bsz):
        table[int(req[b]), : ctx[b]] = perm[cursor: cursor + ctx[b]]
        cursor += ctx[b]
    total = sum(n for _, n in rows)
    q = randn_bf16(gen, total, hq, d)
    start_loc, off = [], 0
    for _, n in rows:
        start_loc.append(off)
        off += n
    cache_lens = [c for c, _ in rows]
    starts = [max(c, c + n - window) if window > 0 else c for c, n in rows]
    v_cache = randn_bf16(gen, num_slots, hkv, d)  # drawn last: the scoring cases' digests do not depend on it
    return dict(q=q, k_cache=k_cache, v_cache=v_cache, slot_table=table, req_indices=req,
                start_loc=torch.tensor(start_loc, dtype=torch.int32), context_lens=torch.tensor(ctx, dtype=torch.int32),
                prompt_cache_lens=torch.tensor(cache_lens, dtype=torch.int32),
                score_starts=torch.tensor(starts, dtype=torch.int32), score_ends=torch.tensor(ctx, dtype=torch.int32),
                max_query_len=max(e - s_ for s_, e in zip(starts, ctx)), max_context_len=width)


def prefill_golden_tokens(inp) -> torch.Tensor:
    """Chunk tokens (indices into q) whose attention output the prefill goldens keep: the first 4 and the last 12
    of every row's chunk (both ends of the causal range), to keep the fixtures small."""
    keep = []
    chunk = (inp["context_lens"] - inp["prompt_cache_lens"]).tolist()
    for s0, n in zip(inp["start_loc"].tolist(), chunk):
        idx = sorted(set(list(range(min(4, n))) + list(range(max(0, n - 12), n))))
        keep += [s0 + i for i in idx]
    return torch.tensor(keep, dtype=torch.long)
