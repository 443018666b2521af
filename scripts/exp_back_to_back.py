"""Repro / experiment: back-to-back fused score launches (eager, per-launch sync, or graph replays) at small
batches (SHAPE=bsz,hq,hkv CTX= BUDGET=); found the stale ring-stage tag bug (tests: test_few_units_per_cta_back_to_back)."""
import os, sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, configs
dev = "cuda:0"
mode = sys.argv[1]
bsz, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '1,32,8').split(','))
ctx, budget = int(os.environ.get('CTX', 32768)), int(os.environ.get('BUDGET', 2048))
gen = torch.Generator(device=dev).manual_seed(1)
max_pages = ctx // 16
n_cols = max_pages - 1
base, prev = configs.quest_page_budget(budget, 16, max_pages)
num_pages = bsz * max_pages
NL = 8
bounds = torch.randn(NL, num_pages, 2, hkv, 128, device=dev, generator=gen, dtype=torch.bfloat16)
pt = torch.randperm(num_pages, device=dev, generator=gen).to(torch.int32).view(bsz, max_pages)
q = torch.randn(bsz, hq, 128, device=dev, generator=gen).to(torch.bfloat16)
ctx_lens = torch.tensor([ctx - 3 * b for b in range(bsz)], dtype=torch.int32, device=dev)
req = torch.arange(bsz, dtype=torch.int32, device=dev)
scores = torch.empty(bsz, ops.score_row_stride(n_cols), dtype=torch.bfloat16, device=dev)
packed = torch.zeros(bsz, (prev + 1) * 16, dtype=torch.int32, device=dev)
lens = torch.zeros(bsz, dtype=torch.int32, device=dev)
tickets = ops.sched_scratch(bsz, dev)
def score(l=0):
    ops.quest_score_select_pack(q, bounds[l], pt, req, ctx_lens, n_cols, hkv, prev, scores, packed, lens, tickets, None)
if mode == "eager8":
    for l in range(NL): score(l)
    torch.cuda.synchronize()
elif mode == "eager8same":
    for l in range(NL): score(0)
    torch.cuda.synchronize()
elif mode == "eager_sync_each":
    for l in range(NL):
        score(l); torch.cuda.synchronize(); print(l, tickets.tolist(), int(lens[0]))
elif mode == "graph1":
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        score(); s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            score(0)
        for _ in range(8): g.replay()
        s.synchronize()
print(mode, "ok", tickets.tolist(), int(lens[0]))
