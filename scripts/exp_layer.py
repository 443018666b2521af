"""Experiment: graph-replayed launch times of the score+select and attention kernels for one shape, across
tuning knobs.  `python scripts/exp_layer.py SHAPE... -- KNOBSET...`, e.g.
    python scripts/exp_layer.py 16,64,8 16,8,1 -- "" "7=1" "7=1,4=2"
SHAPE = bsz,hq,hkv[,ctx[,budget]]; KNOBSET = comma-separated which=value (svb200_set_tuning)."""
import sys

import torch

from exp_common import QuestState, graph_time_us, ops

args = sys.argv[1:]
sep = args.index("--") if "--" in args else len(args)
shapes, knobsets = args[:sep] or ["16,64,8"], args[sep + 1:] or [""]
NL = 8
for shape in shapes:
    f = [int(x) for x in shape.split(",")]
    bsz, hq, hkv = f[:3]
    ctx = f[3] if len(f) > 3 else 131072
    budget = f[4] if len(f) > 4 else 4096
    st = QuestState(bsz, hq, hkv, ctx, budget, layers=NL)
    for ks in knobsets:
        pairs = [tuple(int(v) for v in item.split("=")) for item in ks.split(",") if item]
        for which, value in pairs:
            ops.set_tuning(which, value)
        t_s = graph_time_us(st.score, NL)
        t_a = graph_time_us(st.attn, NL)
        t_b = graph_time_us(st.both, NL)
        t_f = graph_time_us(st.fused, NL) if hq // hkv <= 8 else float("nan")
        print(f"shape {shape:>22s} knobs [{ks:>12s}]  score+select {t_s:7.1f} us ({st.bytes_score() / t_s / 1e3:6.0f} GB/s)  "
              f"attention {t_a:6.1f} us ({st.bytes_attn() / t_a / 1e3:6.0f} GB/s)  layer {t_b:7.1f} us  one launch {t_f:7.1f} us", flush=True)
        for which, _ in pairs:
            ops.set_tuning(which, 0)
    del st
    torch.cuda.empty_cache()
