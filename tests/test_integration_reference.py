"""Boundary proof: INTEGRATION.md's reference-side stubs (integration/sparsevllm_svb200.py) executed against
the REAL reference classes on CPU tensors.  The library is replaced by a recorder that checks every call
against the C prototypes (argument count per include/svb200.h / _native._SIGNATURES), so the claim "drop-in behind
the reference's cache-manager / backend API" is run, not only documented: overridden hooks keep the reference's
signatures, the reference's own page allocator drives the bounds-chunk maps, and each hook reaches the entry
point INTEGRATION.md names.  Needs the reference (build container: /root/reference, else the vendored oracle/_ref)."""
from __future__ import annotations

import importlib.util
import inspect
import os
import sys
from collections import deque
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
PAGE = 16


def _reference_path() -> str | None:
    for cand in ("/root/reference/src", str(ROOT / "oracle" / "_ref")):
        if os.path.exists(os.path.join(cand, "sparsevllm", "__init__.py")):
            return cand
    return None


class _RecorderLib:
    """Stands in for the ctypes library: records calls, checks their arity against the real prototypes."""

    def __init__(self):
        from sparse_vllm_b200 import _native
        self.signatures = _native._SIGNATURES
        self.calls: list[tuple[str, tuple]] = []

    def __getattr__(self, name):
        if name not in self.signatures:
            raise AttributeError(name)
        n_args = len(self.signatures[name][1])

        def fn(*args):
            assert len(args) == n_args, f"{name}: {len(args)} arguments, the prototype has {n_args}"
            for a in args:
                assert a is None or isinstance(a, (int, float)), f"{name}: argument {a!r} is not a plain pointer / number"
            self.calls.append((name, args))
            if name == "svb200_workspace_bytes":
                return 1024
            if name == "svb200_sched_scratch_ints":
                return 1 + 2 * args[0]
            return 0
        return fn

    def names(self):
        return [c[0] for c in self.calls]


@pytest.fixture(scope="module")
def integ():
    ref = _reference_path()
    if ref is None:
        pytest.skip("the reference package is not available on this machine")
    os.environ.setdefault("SPARSEVLLM_PLATFORM", "cpu")
    os.environ.setdefault("CUDA_VISIBLE_DEVICES", os.environ.get("CUDA_VISIBLE_DEVICES", ""))
    if ref not in sys.path:
        sys.path.insert(0, ref)
    spec = importlib.util.spec_from_file_location("sparsevllm_svb200", ROOT / "integration" / "sparsevllm_svb200.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    mod._lib = _RecorderLib()
    return mod


def test_stub_prototypes_match_the_header(integ):
    from sparse_vllm_b200 import _native
    assert integ.ABI_VERSION == _native.ABI_VERSION
    for name, argtypes in integ.ARGTYPES.items():
        assert [t._type_ for t in argtypes] == [t._type_ for t in _native._SIGNATURES[name][1]], name


def test_overridden_hooks_keep_the_reference_signatures(integ):
    from sparsevllm.engine.cache_manager.quest import QuestCacheManager
    from sparsevllm.layers.attention_backend import TritonAttentionBackend
    for sub, base, names in (
            (integ.QuestB200CacheManager, QuestCacheManager,
             ("allocate_kv_cache", "_allocate", "_allocate_batch", "free_seq", "on_kv_stored", "on_forward_end",
              "_build_decode_view_static")),
            (integ.SVB200AttentionBackend, TritonAttentionBackend, ("run_decode",))):
        for name in names:
            a = inspect.signature(getattr(sub, name)).parameters
            b = inspect.signature(getattr(base, name)).parameters
            assert list(a) == list(b), f"{sub.__name__}.{name}: parameters {list(a)} != reference {list(b)}"
            assert [p.kind for p in a.values()] == [p.kind for p in b.values()], f"{sub.__name__}.{name}: parameter kinds"
    # the hooks the generic Attention layer / engine call are still the reference's own methods
    for name in ("build_decode_view", "build_decode_compute_view", "get_layer_kv_cache", "prepare_decode_static"):
        assert getattr(integ.QuestB200CacheManager, name) is getattr(QuestCacheManager, name)


def _manager(integ, *, rows=3, num_pages=96, max_model_len=640, hkv=2, layers=2, token_budget=128):
    from oracle import ref_arms
    m = ref_arms.make_ref_quest_manager(integ.QuestB200CacheManager, device="cpu", layers=layers, hkv=hkv,
                                        num_pages=num_pages, rows=rows, max_model_len=max_model_len,
                                        token_budget=token_budget)
    # what QuestCacheManager.__init__ / allocate_kv_cache set up for the page allocator (quest.py:46-105, 174-194)
    m.free_rows = deque(range(rows))
    m.free_pages_cpu_stack = np.arange(num_pages, dtype=np.int32)[::-1].copy()
    m.free_pages_stack = torch.from_numpy(m.free_pages_cpu_stack.copy())
    m._num_free_pages = num_pages
    m.prefix_cache = None
    m.seq_id_to_cached_pages, m.seq_id_to_prefix_blocks, m.seq_id_to_materialized_blocks = {}, {}, {}
    m.prefix_runtime_states, m.pending_prefix_blocks = {}, {}
    m.metadata_cache = None
    m._init_b200_state()
    return m


def test_reference_allocator_drives_the_bounds_maps_and_hooks_reach_the_library(integ):
    from sparsevllm.utils.context import set_context
    lib = integ._lib
    lib.calls.clear()
    m = _manager(integ)
    hkv, hq = 2, 8
    # ---- prefill through the reference's own _allocate (two chunks for seq 7, one for seq 9)
    set_context(True)
    for seq_id, sizes in ((7, (300, 213)), (9, (256,))):
        for n in sizes:
            slots = m._allocate(seq_id, n)
            assert slots.numel() == n
            m.on_kv_stored(0, torch.zeros(n, hkv, 128, dtype=torch.bfloat16), slots.to(torch.int32))
    assert lib.names() == ["svb200_quest_page_minmax"] * 3
    rows = [m.seq_id_to_row[7], m.seq_id_to_row[9]]
    rec, pt, ut = m.page_rec.numpy(), m.buffer_req_to_page_slots_cpu, m.unit_table_cpu
    for row, n_tok in zip(rows, (513, 256)):
        n_pages = (n_tok + PAGE - 1) // PAGE
        j = np.arange(n_pages)
        assert (ut[row, : (n_pages + PAGE - 1) // PAGE] >= 0).all()
        assert np.array_equal(rec[pt[row, :n_pages]], ut[row, j // PAGE] * PAGE + j % PAGE)
    assert np.array_equal(m.unit_table.numpy(), ut) and (rec >= 0).sum() == 33 + 16
    # ---- a decode step through the reference's _allocate_batch: seq 9 opens logical page 16 -> a new chunk
    set_context(False, is_long_text=True)
    units_before = int((ut[rows[1]] >= 0).sum())
    slots = m._allocate_batch([7, 9], 1)
    assert slots.numel() == 2 and int((ut[rows[1]] >= 0).sum()) == units_before + 1
    assert int(rec[pt[rows[1], 16]]) == int(ut[rows[1], 1]) * PAGE
    # ---- build_decode_view (the reference's own gate, quest.py:1805-1830) -> the provider's one-launch view
    lib.calls.clear()
    m.layer_batch_state.max_context_len = 640
    req = torch.tensor(rows, dtype=torch.int32)
    ctx = torch.tensor([514, 257], dtype=torch.int32)
    m.layer_batch_state.req_indices, m.layer_batch_state.context_lens = req, ctx
    q = torch.zeros(2, hq, 128, dtype=torch.bfloat16)
    packed, req_out, lens = m.build_decode_view(0, q, m.buffer_req_to_token_slots, req, ctx, num_heads=hq, num_kv_heads=hkv)
    assert lib.names() == ["svb200_sched_scratch_ints", "svb200_quest_score_select_pack"]
    args = lib.calls[-1][1]
    assert args[3] == m.num_chunks and args[12:17] == (2, hq, hkv, 39, 7)  # batch, heads, n_cols = 640/16 - 1, prev = 128/16 - 1
    assert tuple(packed.shape) == (2, 8 * PAGE) and req_out.tolist() == [0, 1] and lens.dtype == torch.int32
    # mixed batch -> the two-launch path with the dense table
    lib.calls.clear()
    set_context(False, is_long_text=False)
    packed, _, _ = m.build_decode_view(0, q, m.buffer_req_to_token_slots, req, ctx, num_heads=hq, num_kv_heads=hkv)
    assert lib.names() == ["svb200_quest_page_score", "svb200_quest_select_pack"] and packed.shape[1] == 128
    # ---- on_forward_end and free_seq
    lib.calls.clear()
    m.on_forward_end([SimpleNamespace(seq_id=7), SimpleNamespace(seq_id=9)], False)
    assert lib.names() == ["svb200_quest_decode_page_complete"]
    free_chunks = len(m._free_chunks)
    m.free_seq(9)
    assert len(m._free_chunks) == free_chunks + 2 and (m.unit_table_cpu[rows[1]] == -1).all()
    assert (m.page_rec.numpy() >= 0).sum() == 33


def test_backend_run_decode_reaches_the_attention_entry_point(integ):
    from sparsevllm.engine.cache_manager.base import AttentionViewMeta, DecodeComputeView, ExplicitKVPayload
    lib = integ._lib
    lib.calls.clear()
    be = object.__new__(integ.SVB200AttentionBackend)
    q = torch.zeros(2, 8, 128, dtype=torch.bfloat16)
    kc = torch.zeros(64, 2, 128, dtype=torch.bfloat16)
    score = torch.full((2, 32), -1e20)
    view = DecodeComputeView(
        meta=AttentionViewMeta(active_slots=torch.zeros(2, 32, dtype=torch.int32), req_indices=torch.arange(2, dtype=torch.int32),
                               context_lens=torch.tensor([20, 32], dtype=torch.int32), attn_score=score, max_context_len=32),
        payload=ExplicitKVPayload(k_cache=kc, v_cache=kc.clone()))
    o = be.run_decode(q, view, mid_o=None, mid_o_logexpsum=None, max_len_in_batch=32, block_seq=256, num_heads=8,
                      num_kv_heads=2)
    assert o.shape == q.shape and lib.names() == ["svb200_workspace_bytes", "svb200_sparse_decode_attn"]
    args = lib.calls[-1][1]
    assert args[12] == score.data_ptr() and args[13] == 32 and args[7] == 32 and args[14:17] == (2, 8, 2)
