"""The reference itself as a timed baseline (test / measurement infrastructure, like the rest of oracle/).

Imports the UNMODIFIED reference package from oracle/_ref (vendored by oracle/build_ref.py) and drives its own
Quest code:
* ``RefCpuSample``: `bench.py --impl reference` / `cpu_baseline` — QuestCacheManager.build_decode_view ->
  _build_decode_view_static -> _score_pages_batched (quest.py:1773-1913) on the host cores.  The reference has
  no CPU attention kernel (its decode kernels are Triton), so the attention half of the sample is the torch
  restatement in oracle/attention.py (kind "reference" covers selection, the dominant half; the line says so).
* ``run_incumbent``: `bench.py --impl incumbent` — the reference's GPU path on the same B200: torch selection
  (cuBLAS bmm + ~20 ATen launches per layer) + Triton gqa_flash_decode_stage1 / flash_decode_stage2 with the
  launch configuration its registry resolves on sm_100 (the generic default: BLOCK_SEQ 256, BLOCK_N 16, 2 warps,
  operators/decode_attention.py:106-128; its tuned provider is H100-only).
Nothing under sparse-vllm_b200/ imports this module.
"""
from __future__ import annotations

import os
import sys
import time
from pathlib import Path
from types import SimpleNamespace

REF_ROOT = Path(__file__).resolve().parent / "_ref"
PAGE = 16
HEAD_DIM = 128


def reference_available() -> str | None:
    """None if oracle/_ref holds the vendored package, else why not."""
    if not (REF_ROOT / "sparsevllm" / "__init__.py").exists():
        return "oracle/_ref/sparsevllm is missing (run python oracle/build_ref.py in the build container)"
    return None


def import_reference(platform: str | None):
    if platform is not None:
        os.environ["SPARSEVLLM_PLATFORM"] = platform
    if str(REF_ROOT) not in sys.path:
        sys.path.insert(0, str(REF_ROOT))
    import sparsevllm  # noqa: F401
    from sparsevllm.engine.cache_manager.quest import QuestCacheManager
    return QuestCacheManager


def make_ref_quest_manager(QuestCacheManager, *, device, layers, hkv, num_pages, rows, max_model_len, token_budget,
                           skip_layers=0, alloc_kv=True):
    """The reference tests' ``object.__new__`` fixture (tests/test_prefix_cache.py:208-235), as in gen_golden.py."""
    import numpy as np
    import torch
    m = object.__new__(QuestCacheManager)
    m.config = SimpleNamespace(quest_skip_layers=skip_layers, quest_token_budget=token_budget)
    m.runtime_layout = SimpleNamespace(kv_layer_index=lambda layer_idx: int(layer_idx) % layers)
    m.device = torch.device(device)
    m.enable_prefix_caching = False
    m.page_size = PAGE
    m.num_pages = num_pages
    m.num_kv_heads = hkv
    m.head_dim = HEAD_DIM
    m.num_kv_layers = layers
    m.max_model_len = max_model_len
    m.max_pages_per_row = (max_model_len + PAGE - 1) // PAGE
    m.page_offsets_i32 = torch.arange(PAGE, dtype=torch.int32, device=device)
    m.page_offsets_i64 = m.page_offsets_i32.to(torch.int64)
    m.kv_cache = (torch.zeros(2, layers, num_pages * PAGE, hkv, HEAD_DIM, dtype=torch.bfloat16, device=device)
                  if alloc_kv else None)
    m.metadata_cache = torch.zeros(2, layers, num_pages, hkv, HEAD_DIM, dtype=torch.bfloat16, device=device)
    m.buffer_req_to_token_slots = torch.zeros((rows, max_model_len), dtype=torch.int32, device=device)
    m.buffer_req_to_page_slots = torch.full((rows, m.max_pages_per_row), -1, dtype=torch.int32, device=device)
    m.buffer_req_to_page_slots_cpu = m.buffer_req_to_page_slots.cpu().numpy()
    m.layer_batch_state = SimpleNamespace(max_context_len=None)
    m.row_seq_lens = np.zeros((rows,), dtype=np.int32)
    m.seq_id_to_row = {}
    m._prefill_metadata_full_pages = False
    m.prefix_offload_controller = None
    return m


def _fill_random_state(m, rows, ctx, seed, device):
    """Random bounds / KV and a shuffled page assignment (what timing needs; parity lives in tests/)."""
    import torch
    g = torch.Generator(device=device).manual_seed(seed)
    pages_per_row = (ctx + PAGE - 1) // PAGE
    perm = torch.randperm(m.num_pages, generator=g, device=device).to(torch.int32)
    m.buffer_req_to_page_slots[:, :pages_per_row] = perm[: rows * pages_per_row].view(rows, pages_per_row)
    slots = (m.buffer_req_to_page_slots[:, :pages_per_row, None] * PAGE + m.page_offsets_i32[None, None, :]).reshape(rows, -1)
    m.buffer_req_to_token_slots[:, : slots.shape[1]] = slots[:, : m.max_model_len]
    for layer in range(m.num_kv_layers):  # layer by layer: a whole-tensor randn would double the footprint
        mx = torch.randn(m.num_pages, m.num_kv_heads, HEAD_DIM, generator=g, device=device, dtype=torch.bfloat16)
        m.metadata_cache[0, layer] = mx
        m.metadata_cache[1, layer] = mx - torch.rand(mx.shape, generator=g, device=device, dtype=torch.bfloat16)
        if m.kv_cache is not None:
            m.kv_cache[0, layer].normal_(generator=g)
            m.kv_cache[1, layer].normal_(generator=g)


class RefCpuSample:
    """`rows` sequences x ONE layer of the workload on the host cores through the reference's own view builder."""

    def __init__(self, w: dict, rows: int):
        import torch
        QuestCacheManager = import_reference("cpu")
        from sparsevllm.utils.context import set_context
        from oracle import attention as oattn
        self.torch, self.oattn, self.w, self.rows = torch, oattn, w, rows
        torch.set_num_threads(os.cpu_count() or 1)
        self.cores = torch.get_num_threads()
        hkv, hq, ctx = w["hkv"], w["hq"], w["ctx"]
        pages_per_row = (ctx + PAGE - 1) // PAGE
        self.m = make_ref_quest_manager(QuestCacheManager, device="cpu", layers=1, hkv=hkv, num_pages=rows * pages_per_row,
                                        rows=rows, max_model_len=ctx, token_budget=w["budget"],
                                        skip_layers=0)
        _fill_random_state(self.m, rows, ctx, seed=5, device="cpu")
        self.m.layer_batch_state.max_context_len = ctx
        set_context(False, is_long_text=True)
        g = torch.Generator().manual_seed(6)
        self.q = torch.randn(rows, hq, HEAD_DIM, generator=g).to(torch.bfloat16)
        self.ctx_lens = torch.tensor([ctx - 3 - b for b in range(rows)], dtype=torch.int32)
        self.req = torch.arange(rows, dtype=torch.int32)

    def run_once(self) -> float:
        w, m = self.w, self.m
        t0 = time.perf_counter()
        packed, req, lens = m.build_decode_view(0, self.q, m.buffer_req_to_token_slots, self.req, self.ctx_lens,
                                                num_heads=w["hq"], num_kv_heads=w["hkv"])
        self.oattn.gqa_decode(self.q, m.kv_cache[0, 0], m.kv_cache[1, 0], packed, req, lens,
                              max_len_in_batch=packed.shape[1])
        return time.perf_counter() - t0

    def tokens_per_s(self, seconds: float) -> float:
        return self.rows / (seconds * self.w["layers"])

    def step_ms(self, seconds: float) -> float:
        """One full decode step (all rows, all layers) at the sample's rate."""
        return seconds * 1e3 * self.w["layers"] * self.w["bsz"] / self.rows

    def describe(self) -> str:
        w = self.w
        return (f"{self.rows} sequence(s) x 1 layer of {w['key']} (ctx {w['ctx']}, {w['hq']}q/{w['hkv']}kv, budget {w['budget']}): "
                "the UNMODIFIED reference's QuestCacheManager.build_decode_view (page gather, 2 bf16 bmm, mask, top-k, pack; "
                "oracle/_ref) + torch-restated sparse attention (the reference has no CPU attention kernel), scaled by "
                "rows/(t*layers)")


def run_incumbent(w: dict, steps: int, warmup: int, phys_layers: int = 4) -> dict:
    """The reference's own GPU path on this box: per layer build_decode_view (torch) + Triton stage 1 + stage 2."""
    import torch
    QuestCacheManager = import_reference(None)
    from sparsevllm.kernels.triton.flash_decoding_stage2 import flash_decode_stage2
    from sparsevllm.kernels.triton.gqa_flash_decoding_stage1 import flash_decode_stage1
    from sparsevllm.utils.context import set_context
    dev = torch.device("cuda", torch.cuda.current_device())
    L, B, hq, hkv, ctx = w["layers"], w["bsz"], w["hq"], w["hkv"], w["ctx"]
    pages_per_row = (ctx + PAGE - 1) // PAGE
    m = make_ref_quest_manager(QuestCacheManager, device=dev, layers=phys_layers, hkv=hkv, num_pages=B * pages_per_row,
                               rows=B, max_model_len=ctx, token_budget=w["budget"])
    _fill_random_state(m, B, ctx, seed=7, device=dev)
    m.layer_batch_state.max_context_len = ctx
    set_context(False, is_long_text=True)
    g = torch.Generator(device=dev).manual_seed(8)
    q = torch.randn(L, B, hq, HEAD_DIM, generator=g, device=dev, dtype=torch.bfloat16)
    ctx_lens = torch.tensor([ctx - 3 - (b * 7) % PAGE for b in range(B)], dtype=torch.int32, device=dev)
    req = torch.arange(B, dtype=torch.int32, device=dev)
    out = torch.empty(L, B, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16)
    block_seq = 256
    prev = max(3, w["budget"] // PAGE)
    width = prev * PAGE
    nblk = (width + block_seq - 1) // block_seq
    mid_o = torch.empty(B, hq, nblk, HEAD_DIM, dtype=torch.float32, device=dev)
    mid_lse = torch.empty(B, hq, nblk, dtype=torch.float32, device=dev)
    sel_ev, attn_ev = [], []

    def step(timed: bool):
        for layer in range(L):
            e = [torch.cuda.Event(enable_timing=True) for _ in range(3)] if timed else None
            if timed:
                e[0].record()
            packed, rq, lens = m.build_decode_view(layer, q[layer], m.buffer_req_to_token_slots, req, ctx_lens,
                                                   num_heads=hq, num_kv_heads=hkv)
            if timed:
                e[1].record()
            k_cache, v_cache = m.kv_cache[0, layer % phys_layers], m.kv_cache[1, layer % phys_layers]
            flash_decode_stage1(q[layer], k_cache, v_cache, packed, rq, lens, int(packed.shape[1]), mid_o, mid_lse, block_seq)
            flash_decode_stage2(mid_o, mid_lse, lens, out[layer], block_seq)
            if timed:
                e[2].record()
                sel_ev.append((e[0], e[1]))
                attn_ev.append((e[1], e[2]))

    for _ in range(max(warmup, 3)):
        step(False)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step(False)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    step(True)
    torch.cuda.synchronize()
    sel_ms = sum(a.elapsed_time(b) for a, b in sel_ev) / len(sel_ev)
    attn_ms = sum(a.elapsed_time(b) for a, b in attn_ev) / len(attn_ev)
    return {"impl": "incumbent", "what": "UNMODIFIED reference on this GPU: QuestCacheManager.build_decode_view (torch/cuBLAS) + "
            "Triton gqa_flash_decode_stage1 + flash_decode_stage2, eager launches (its decode graph captures the same "
            "kernels), generic sm_100 launch config (BLOCK_SEQ 256, BLOCK_N 16, 2 warps)",
            "workload": w["name"], "value": B / (ms * 1e-3), "unit": "tok/s", "ms_per_step": ms, "steps": steps,
            "select_ms_per_layer": sel_ms, "attention_ms_per_layer": attn_ms, "physical_kv_pools": phys_layers}
