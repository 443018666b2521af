"""One launch per sparse decode layer (svb200_quest_decode_layer) against the two launches it replaces
(svb200_quest_(store_)score_select_pack, then svb200_sparse_decode_attn), which the other GPU tests pin on the
reference's goldens: scores, selected pages, packed slots and lengths bit for bit (same code, and selection does not
depend on who scored what); the attention output within the attention tolerance (the CTA cut of the fused launch
differs, so fp32 partial sums are merged in another order) and against the oracle's attention."""
from __future__ import annotations

import pytest
import torch

from helpers import packed_bounds
from oracle import attention as oattn
from oracle.cases import chunk_maps  # index helper only

pytestmark = pytest.mark.gpu
PAGE = 16
DEV = "cuda:0"


class _Layer:
    def __init__(self, bsz, hq, hkv, ctx, budget, seed, ragged=True):
        from sparse_vllm_b200 import configs, ops
        gen = torch.Generator(device=DEV).manual_seed(seed)
        self.bsz, self.hq, self.hkv = bsz, hq, hkv
        max_pages = ctx // PAGE
        self.n_cols = max_pages - 1
        _, self.prev = configs.quest_page_budget(budget, PAGE, max_pages)
        num_pages = bsz * max_pages
        self.num_slots = num_pages * PAGE
        bmax = torch.randn(num_pages, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        bmin = bmax - torch.rand(num_pages, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        self.pt = torch.randperm(num_pages, device=DEV, generator=gen).to(torch.int32).view(bsz, max_pages)
        ut, rec, num_chunks = chunk_maps(self.pt, num_pages, seed=seed)
        self.ut, self.rec = ut.to(DEV), rec.to(DEV)
        self.bounds = packed_bounds(torch.stack((bmax, bmin), dim=1), self.rec, num_chunks, DEV)
        self.kc = torch.randn(self.num_slots, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        self.vc = torch.randn(self.num_slots, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        self.q = torch.randn(bsz, hq, 128, device=DEV, generator=gen).to(torch.bfloat16)
        step = 37 if ragged else 0
        self.ctx_lens = torch.tensor([ctx - step * b - 3 for b in range(bsz)], dtype=torch.int32, device=DEV)
        self.req = torch.arange(bsz, dtype=torch.int32, device=DEV)
        self.k_new = torch.randn(bsz, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        self.v_new = torch.randn(bsz, hkv, 128, device=DEV, generator=gen).to(torch.bfloat16)
        # the step's token of row b lands in the row's last page
        pos = self.ctx_lens.long() - 1
        self.slot_mapping = (self.pt[torch.arange(bsz, device=DEV), pos // PAGE].long() * PAGE + pos % PAGE).to(torch.int32)
        self.ops = ops

    def buffers(self):
        ops = self.ops
        return dict(scores=torch.zeros(self.bsz, ops.score_row_stride(self.n_cols), dtype=torch.bfloat16, device=DEV),
                    packed=torch.zeros(self.bsz, (self.prev + 1) * PAGE, dtype=torch.int32, device=DEV),
                    lens=torch.zeros(self.bsz, dtype=torch.int32, device=DEV),
                    sel=torch.zeros(self.bsz, self.prev + 1, dtype=torch.int32, device=DEV),
                    out=torch.zeros(self.bsz, self.hq, 128, dtype=torch.bfloat16, device=DEV),
                    sched=ops.sched_scratch(self.bsz, DEV), ws=ops.attn_workspace(self.bsz, self.hq, self.hkv, DEV))

    def store(self, kc, vc):
        return self.ops.DecodeStore(self.k_new, self.v_new, kc, vc, self.slot_mapping, True, self.rec)

    def two_launches(self, b, kc, vc, bounds, with_store):
        ops = self.ops
        ops.quest_score_select_pack(self.q, bounds, self.ut, self.pt, self.req, self.ctx_lens, self.n_cols, self.hkv,
                                    self.prev, b["scores"], b["packed"], b["lens"], b["sched"], b["sel"],
                                    store=self.store(kc, vc) if with_store else None)
        ops.sparse_decode_attn(self.q, kc, vc, b["packed"], self.req, b["lens"], out=b["out"], workspace=b["ws"])

    def one_launch(self, b, kc, vc, bounds, with_store):
        self.ops.quest_decode_layer(self.q, bounds, self.ut, self.pt, self.req, self.ctx_lens, self.n_cols, self.hkv,
                                    self.prev, b["scores"], b["packed"], b["lens"], b["sched"], kc, vc, out=b["out"],
                                    workspace=b["ws"], selected_pages=b["sel"],
                                    store=self.store(kc, vc) if with_store else None)


def _compare(layer, want, got, tag):
    n = layer.n_cols
    assert torch.equal(got["scores"][:, :n].view(torch.int16), want["scores"][:, :n].view(torch.int16)), f"{tag}: scores"
    for key in ("packed", "lens", "sel"):
        assert torch.equal(got[key], want[key]), f"{tag}: {key}"
    assert int(want["lens"].min()) > 0
    torch.testing.assert_close(got["out"].float(), want["out"].float(), atol=2e-2, rtol=1e-2)
    assert not bool(got["sched"].any()), f"{tag}: the scheduler scratch is not left zeroed"


@pytest.mark.parametrize("with_store", [False, True])
@pytest.mark.parametrize("bsz,hq,hkv,ctx,budget", [(5, 28, 4, 16384, 512), (3, 64, 8, 8192, 256), (4, 8, 1, 8192, 256),
                                                   (2, 16, 2, 4096, 128), (1, 32, 8, 32768, 2048), (16, 8, 1, 16384, 1024)])
def test_one_launch_layer_matches_the_two_launches(bsz, hq, hkv, ctx, budget, with_store):
    layer = _Layer(bsz, hq, hkv, ctx, budget, seed=bsz * 100 + hkv)
    want, got = layer.buffers(), layer.buffers()
    kc_a, vc_a, bd_a = layer.kc.clone(), layer.vc.clone(), layer.bounds.clone()
    kc_b, vc_b, bd_b = layer.kc.clone(), layer.vc.clone(), layer.bounds.clone()
    layer.two_launches(want, kc_a, vc_a, bd_a, with_store)
    torch.cuda.synchronize()
    for rep in range(3):  # back to back on the same scratch / workspace: every launch leaves them reusable
        got["out"].zero_()
        layer.one_launch(got, kc_b, vc_b, bd_b, with_store)
        torch.cuda.synchronize()
        _compare(layer, want, got, f"launch {rep}")
    if with_store:
        assert torch.equal(kc_a, kc_b) and torch.equal(vc_a, vc_b) and torch.equal(bd_a, bd_b)
    # and against the oracle's attention over the packed slots
    ref = oattn.gqa_decode(layer.q.cpu(), kc_b.cpu(), vc_b.cpu(), got["packed"].cpu(),
                           torch.arange(bsz, dtype=torch.int32), got["lens"].cpu(), max_len_in_batch=got["packed"].shape[1])
    torch.testing.assert_close(got["out"].cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)


@pytest.mark.parametrize("max_ctas", [1, 3, 40])
def test_one_launch_layer_with_few_ctas(max_ctas):
    from sparse_vllm_b200 import ops
    layer = _Layer(3, 28, 4, 8192, 256, seed=5)
    want, got = layer.buffers(), layer.buffers()
    layer.two_launches(want, layer.kc, layer.vc, layer.bounds, False)
    torch.cuda.synchronize()
    ops.set_max_ctas(max_ctas)
    try:
        layer.one_launch(got, layer.kc, layer.vc, layer.bounds, False)
        torch.cuda.synchronize()
    finally:
        ops.set_max_ctas(0)
    _compare(layer, want, got, f"max_ctas {max_ctas}")


def test_one_launch_layer_replays_in_a_cuda_graph():
    layer = _Layer(4, 32, 8, 8192, 512, seed=9)
    want, got = layer.buffers(), layer.buffers()
    layer.two_launches(want, layer.kc, layer.vc, layer.bounds, False)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        layer.one_launch(got, layer.kc, layer.vc, layer.bounds, False)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(4):
                layer.one_launch(got, layer.kc, layer.vc, layer.bounds, False)
        for _ in range(3):
            got["out"].zero_()
            g.replay()
        s.synchronize()
    torch.cuda.synchronize()
    _compare(layer, want, got, "graph replay")
