"""svb200_quest_score_select_pack_static: the fused score + select + pack launch for callers that cannot allocate
device memory (its scheduler scratch is a static device array inside the library).  No counterpart in the
reference.  Collected last on purpose (file name):
a failure here must not hide the parity tests.

The GPU test calls the new entry point through the C ABI twice in a row (different inputs and batch sizes, so the
second call finds the array as the first one left it) and compares every output with the existing entry point's,
bit for bit: both run the same kernel and its results do not depend on which CTA scored what."""
from __future__ import annotations

import pytest
import torch

from helpers import packed_bounds
from oracle.cases import chunk_maps  # index helper only

PAGE = 16
DEV = "cuda:0"


def _case(bsz, hq, hkv, ctx, budget, seed):
    from sparse_vllm_b200 import configs, ops
    gen = torch.Generator(device=DEV).manual_seed(seed)
    max_pages = ctx // PAGE
    n_cols = max_pages - 1
    _, prev = configs.quest_page_budget(budget, PAGE, max_pages)
    num_pages = bsz * max_pages
    bmax = torch.randn(num_pages, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
    bmin = bmax - torch.rand(num_pages, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
    pt = torch.randperm(num_pages, device=DEV, generator=gen).to(torch.int32).view(bsz, max_pages)
    ut, rec, num_chunks = chunk_maps(pt, num_pages, seed=seed)
    ut, rec = ut.to(DEV), rec.to(DEV)
    bounds = packed_bounds(torch.stack((bmax, bmin), dim=1), rec, num_chunks, DEV)
    q = torch.randn(bsz, hq, 128, device=DEV, generator=gen).to(torch.bfloat16)
    ctx_lens = torch.tensor([ctx - 5 * b for b in range(bsz)], dtype=torch.int32, device=DEV)
    req = torch.arange(bsz, dtype=torch.int32, device=DEV)

    def outputs():
        return dict(scores=torch.zeros(bsz, ops.score_row_stride(n_cols), dtype=torch.bfloat16, device=DEV),
                    packed=torch.zeros(bsz, (prev + 1) * PAGE, dtype=torch.int32, device=DEV),
                    lens=torch.zeros(bsz, dtype=torch.int32, device=DEV),
                    sel=torch.zeros(bsz, prev + 1, dtype=torch.int32, device=DEV))
    return dict(q=q, bounds=bounds, ut=ut, pt=pt, req=req, ctx_lens=ctx_lens, n_cols=n_cols, hkv=hkv, prev=prev,
                bsz=bsz, outputs=outputs)


@pytest.mark.gpu
def test_static_scratch_entry_matches_the_caller_scratch_entry_twice_in_a_row():
    from sparse_vllm_b200 import ops
    # two calls in a row on the shared array: another shape, batch and data the second time
    cases = [_case(5, 28, 4, 16384, 512, seed=7), _case(3, 64, 8, 8192, 256, seed=8)]
    for call, c in enumerate(cases):
        want, got = c["outputs"](), c["outputs"]()
        ops.quest_score_select_pack(c["q"], c["bounds"], c["ut"], c["pt"], c["req"], c["ctx_lens"], c["n_cols"],
                                    c["hkv"], c["prev"], want["scores"], want["packed"], want["lens"],
                                    ops.sched_scratch(c["bsz"], DEV), want["sel"])
        torch.cuda.synchronize()
        free_before = torch.cuda.mem_get_info()[0]
        torch_before = torch.cuda.memory_allocated()
        ops.quest_score_select_pack_static(c["q"], c["bounds"], c["ut"], c["pt"], c["req"], c["ctx_lens"], c["n_cols"],
                                           c["hkv"], c["prev"], got["scores"], got["packed"], got["lens"], got["sel"])
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == torch_before, "the static entry point's wrapper allocated through torch"
        if call > 0:  # (the first call of a process may still load the module's code lazily)
            assert torch.cuda.mem_get_info()[0] >= free_before, "device memory was allocated during the call"
        n = c["n_cols"]
        assert torch.equal(got["scores"][:, :n].view(torch.int16), want["scores"][:, :n].view(torch.int16)), f"call {call}: scores"
        for key in ("packed", "lens", "sel"):
            assert torch.equal(got[key], want[key]), f"call {call}: {key}"
        assert int(want["lens"].min()) > 0  # the comparison is not of two untouched buffers


def test_static_scratch_entry_rejects_batches_beyond_its_maximum(native_lib):
    """Validation precedes every CUDA call, so this needs no GPU."""
    from sparse_vllm_b200 import ops
    too_many = ops.STATIC_MAX_BATCH + 1
    rc = native_lib.svb200_quest_score_select_pack_static(None, 0, None, 0, None, 0, None, 0, None, None, None, 0,
                                                          too_many, 8, 1, 127, 15, None, 0, None, None, None)
    assert rc == 1
    msg = native_lib.svb200_last_error()
    assert b"SVB200_STATIC_MAX_BATCH" in msg and str(too_many).encode() in msg
