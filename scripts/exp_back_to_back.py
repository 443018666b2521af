"""Repro / experiment: back-to-back fused score launches (eager, per-launch sync, or graph replays) at small
batches (SHAPE=bsz,hq,hkv CTX= BUDGET=); found the stale ring-stage tag bug (tests: test_few_units_per_cta_back_to_back)."""
import os
import sys

import torch

from exp_common import QuestState

mode = sys.argv[1]
bsz, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '1,32,8').split(','))
NL = 8
st = QuestState(bsz, hq, hkv, int(os.environ.get('CTX', 32768)), int(os.environ.get('BUDGET', 2048)), layers=NL, kv=False)
score, tickets, lens = st.score, st.tickets, st.lens
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
