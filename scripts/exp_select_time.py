"""Experiment: duration of the standalone top-k select + pack kernel (one CTA per row)."""
import sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, configs
dev = "cuda:0"
for bsz, ctx, budget in ((32, 131072, 4096), (1, 131072, 4096), (32, 32768, 2048)):
    max_pages = ctx // 16
    n_cols = max_pages - 1
    base, prev = configs.quest_page_budget(budget, 16, max_pages)
    scores = torch.randn(bsz, ops.score_row_stride(n_cols), device=dev).to(torch.bfloat16)
    pt = torch.randperm(bsz * max_pages, device=dev).to(torch.int32).view(bsz, max_pages)
    ctx_lens = torch.full((bsz,), ctx - 3, dtype=torch.int32, device=dev)
    req = torch.arange(bsz, dtype=torch.int32, device=dev)
    packed = torch.zeros(bsz, (prev + 1) * 16, dtype=torch.int32, device=dev)
    lens = torch.zeros(bsz, dtype=torch.int32, device=dev)
    for _ in range(3):
        ops.quest_select_pack(scores, n_cols, pt, req, ctx_lens, prev, packed, lens)
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        ops.quest_select_pack(scores, n_cols, pt, req, ctx_lens, prev, packed, lens)
        stream.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=stream):
            for _ in range(50):
                ops.quest_select_pack(scores, n_cols, pt, req, ctx_lens, prev, packed, lens)
        g.replay(); stream.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        g.replay()
        e1.record(stream); stream.synchronize()
    print(f"select_pack bsz {bsz} n_cols {n_cols} k {prev}: {e0.elapsed_time(e1) / 50 * 1e3:.2f} us per launch (graph of 50, L2-warm)")
