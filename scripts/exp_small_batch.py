"""Experiment: small-batch (cfg2: bs 1, ctx 32K, budget 2048) launch times vs the CTA cap, graph-replayed."""
import os, sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, configs
dev = "cuda:0"
bsz, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '1,32,8').split(','))
ctx, budget = int(os.environ.get('CTX', 32768)), int(os.environ.get('BUDGET', 2048))
gen = torch.Generator(device=dev).manual_seed(1)
max_pages = ctx // 16
n_cols = max_pages - 1
base, prev = configs.quest_page_budget(budget, 16, max_pages)
num_pages = bsz * max_pages
NL = 8  # distinct layers so that L2 does not help
bounds = torch.randn(NL, num_pages, 2, hkv, 128, device=dev, generator=gen, dtype=torch.bfloat16)
kc = torch.randn(NL, num_pages * 16, hkv, 128, device=dev, generator=gen, dtype=torch.bfloat16)
vc = torch.randn(NL, num_pages * 16, hkv, 128, device=dev, generator=gen, dtype=torch.bfloat16)
pt = torch.randperm(num_pages, device=dev, generator=gen).to(torch.int32).view(bsz, max_pages)
q = torch.randn(bsz, hq, 128, device=dev, generator=gen).to(torch.bfloat16)
ctx_lens = torch.tensor([ctx - 3 * b for b in range(bsz)], dtype=torch.int32, device=dev)
req = torch.arange(bsz, dtype=torch.int32, device=dev)
scores = torch.empty(bsz, ops.score_row_stride(n_cols), dtype=torch.bfloat16, device=dev)
packed = torch.zeros(bsz, (prev + 1) * 16, dtype=torch.int32, device=dev)
lens = torch.zeros(bsz, dtype=torch.int32, device=dev)
tickets = ops.sched_scratch(bsz, dev)
out = torch.empty_like(q)
ws = ops.attn_workspace(bsz, hq, hkv, dev)
hpc = int(os.environ.get('HPC', 4))
ops.set_tuning(4, int(os.environ.get('COPY', '0')))  # 1: LDGSTS instead of bulk copies in the score kernel


def timed(fn, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        fn(); s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for l in range(NL):
                fn(l)
        g.replay(); s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s); s.synchronize()
    return e0.elapsed_time(e1) * 1e3 / (reps * NL)


def score(l=0):
    ops.quest_score_select_pack(q, bounds[l], pt, req, ctx_lens, n_cols, hkv, prev, scores, packed, lens, tickets, None)


def attn(l=0):
    ops.sparse_decode_attn(q, kc[l], vc[l], packed, req, lens, out=out, workspace=ws, heads_per_cta=hpc)


for cap in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0,16,32,64,96").split(",")]:
    ops.set_max_ctas(cap)
    which = os.environ.get('WHICH', 'both')
    if which in ('both', 'score'):
        print(f"cap {cap:4d}: score+select {timed(score):6.1f} us", flush=True)
    if which in ('both', 'attn'):
        print(f"cap {cap:4d}: attention    {timed(attn):6.1f} us", flush=True)
ops.set_max_ctas(0)
if os.environ.get('MINKB'):
    for kb in [int(x) for x in os.environ['MINKB'].split(',')]:
        ops.set_tuning(1, kb)
        print(f"min KB per CTA {kb:4d}: attention {timed(attn):6.1f} us", flush=True)
    ops.set_tuning(1, 0)
