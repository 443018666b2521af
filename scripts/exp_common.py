"""Shared set-up of the experiment scripts: random Quest state of a given shape on cuda:0 in the provider's layouts."""
import os
import sys

import torch

sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import svb200_import  # noqa: E402,F401
from oracle.cases import chunk_maps  # noqa: E402  (index helper only)
from sparse_vllm_b200 import configs, ops  # noqa: E402

DEV = "cuda:0"


class QuestState:
    """`layers` distinct layers (so that L2 does not help) of bounds chunks + KV for bsz rows of ~ctx tokens."""

    def __init__(self, bsz, hq, hkv, ctx, budget, layers=1, kv=True, seed=1):
        gen = torch.Generator(device=DEV).manual_seed(seed)
        self.bsz, self.hq, self.hkv, self.ctx, self.budget, self.layers = bsz, hq, hkv, ctx, budget, layers
        self.max_pages = ctx // 16
        self.n_cols = self.max_pages - 1
        _, self.prev = configs.quest_page_budget(budget, 16, self.max_pages)
        self.num_pages = bsz * self.max_pages
        self.pt = torch.randperm(self.num_pages, device=DEV, generator=gen).to(torch.int32).view(bsz, self.max_pages)
        ut, rec, self.num_chunks = chunk_maps(self.pt, self.num_pages, seed=seed)
        self.ut, self.rec = ut.to(DEV), rec.to(DEV)
        self.bounds = torch.randn((layers, *ops.bounds_shape(self.num_chunks, hkv)), device=DEV, generator=gen,
                                  dtype=torch.bfloat16)  # random bytes are as good as packed ones for timing
        if kv:
            self.kc = torch.randn(layers, self.num_pages * 16, hkv, 128, device=DEV, generator=gen, dtype=torch.bfloat16)
            self.vc = torch.randn(layers, self.num_pages * 16, hkv, 128, device=DEV, generator=gen, dtype=torch.bfloat16)
        self.q = torch.randn(bsz, hq, 128, device=DEV, generator=gen).to(torch.bfloat16)
        self.ctx_lens = torch.tensor([ctx - 3 * b for b in range(bsz)], dtype=torch.int32, device=DEV)
        self.req = torch.arange(bsz, dtype=torch.int32, device=DEV)
        self.scores = torch.empty(bsz, ops.score_row_stride(self.n_cols), dtype=torch.bfloat16, device=DEV)
        self.packed = torch.zeros(bsz, (self.prev + 1) * 16, dtype=torch.int32, device=DEV)
        self.lens = torch.zeros(bsz, dtype=torch.int32, device=DEV)
        self.tickets = ops.sched_scratch(bsz, DEV)
        self.out = torch.empty_like(self.q)
        self.ws = ops.attn_workspace(bsz, hq, hkv, DEV)

    def score(self, layer=0):
        ops.quest_score_select_pack(self.q, self.bounds[layer % self.layers], self.ut, self.pt, self.req, self.ctx_lens,
                                    self.n_cols, self.hkv, self.prev, self.scores, self.packed, self.lens, self.tickets, None)

    def attn(self, layer=0, hpc=0):
        ops.sparse_decode_attn(self.q, self.kc[layer % self.layers], self.vc[layer % self.layers], self.packed, self.req,
                               self.lens, out=self.out, workspace=self.ws, heads_per_cta=hpc)

    def fused(self, layer=0):
        ops.quest_decode_layer(self.q, self.bounds[layer % self.layers], self.ut, self.pt, self.req, self.ctx_lens,
                               self.n_cols, self.hkv, self.prev, self.scores, self.packed, self.lens, self.tickets,
                               self.kc[layer % self.layers], self.vc[layer % self.layers], out=self.out, workspace=self.ws)

    def both(self, layer=0):
        self.score(layer)
        self.attn(layer)

    def bytes_score(self):
        return self.bsz * (2 * self.n_cols * self.hkv * 128 * 2 + self.hq * 256 + 4 * self.max_pages + 4 * (self.prev + 1) * 16)

    def bytes_attn(self):
        t = (self.prev + 1) * 16
        return self.bsz * (2 * t * self.hkv * 128 * 2 + 2 * self.hq * 256 + 4 * t)


def graph_time_us(fn, n_layers, reps=5):
    """fn(layer) captured for n_layers layers in one CUDA graph; average microseconds per fn call."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(0)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for layer in range(n_layers):
                fn(layer)
        g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s)
        s.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * n_layers)
