"""Experiment: does the ORDER of the launches matter?  One CUDA graph with 8 score+select launches followed by 8 attention
launches against one with the same 16 launches interleaved (score, attention, score, ...), same data.
SHAPE=bsz,hq,hkv python scripts/exp_boundary.py"""
import os

import torch

from exp_common import QuestState

f = [int(x) for x in os.environ.get("SHAPE", "16,64,8").split(",")]
st = QuestState(f[0], f[1], f[2], f[3] if len(f) > 3 else 131072, f[4] if len(f) > 4 else 4096, layers=8)
NL = 8


def time_graph(seq, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for kind, layer in seq[:2]:
            (st.score if kind == "s" else st.attn)(layer)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for kind, layer in seq:
                (st.score if kind == "s" else st.attn)(layer)
        g.replay()
        s.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(s)
        for _ in range(reps):
            g.replay()
        e1.record(s)
        s.synchronize()
        return e0.elapsed_time(e1) * 1e3 / reps


blocked = [("s", l) for l in range(NL)] + [("a", l) for l in range(NL)]
inter = [x for l in range(NL) for x in (("s", l), ("a", l))]
pairs = [x for l in range(0, NL, 2) for x in (("s", l), ("s", l + 1), ("a", l), ("a", l + 1))]
only_s = [("s", l) for l in range(NL)]
only_a = [("a", l) for l in range(NL)]
for name, seq in (("8 x score", only_s), ("8 x attention", only_a), ("8 x score then 8 x attention", blocked),
                  ("score, attention interleaved", inter), ("2 score, 2 attention, ...", pairs)):
    print(f"{name:34s} {time_graph(seq):9.1f} us per replay", flush=True)
