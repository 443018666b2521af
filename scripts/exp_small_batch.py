"""Experiment: small launches (cfg2: bs 1, ctx 32K, budget 2048; TP shards of cfg5) vs the CTA cap, graph-replayed.
SHAPE=bsz,hq,hkv CTX=.. BUDGET=.. WHICH=both|score|attn python scripts/exp_small_batch.py 0,16,32,64"""
import os
import sys

from exp_common import QuestState, graph_time_us, ops

bsz, hq, hkv = (int(x) for x in os.environ.get("SHAPE", "1,32,8").split(","))
ctx, budget = int(os.environ.get("CTX", 32768)), int(os.environ.get("BUDGET", 2048))
NL = 8  # distinct layers so that L2 does not help
st = QuestState(bsz, hq, hkv, ctx, budget, layers=NL)
hpc = int(os.environ.get("HPC", 0))
which = os.environ.get("WHICH", "both")
for cap in [int(x) for x in (sys.argv[1] if len(sys.argv) > 1 else "0,16,32,64,96").split(",")]:
    ops.set_max_ctas(cap)
    if which in ("both", "score"):
        print(f"cap {cap:4d}: score+select {graph_time_us(st.score, NL):6.1f} us", flush=True)
    if which in ("both", "attn"):
        print(f"cap {cap:4d}: attention    {graph_time_us(lambda l: st.attn(l, hpc), NL):6.1f} us", flush=True)
ops.set_max_ctas(0)
if os.environ.get("MINKB"):
    for kb in [int(x) for x in os.environ["MINKB"].split(",")]:
        ops.set_tuning(1, kb)
        print(f"min KB per CTA {kb:4d}: attention {graph_time_us(lambda l: st.attn(l, hpc), NL):6.1f} us", flush=True)
    ops.set_tuning(1, 0)
