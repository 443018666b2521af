"""Eager launches of the score+select and attention kernels of one shape (for ncu captures).
SHAPE=bsz,hq,hkv[,ctx[,budget]] LAYERS=n python scripts/exp_ncu_shape.py"""
import os

import torch

from exp_common import QuestState

f = [int(x) for x in os.environ.get("SHAPE", "16,8,1").split(",")]
st = QuestState(f[0], f[1], f[2], f[3] if len(f) > 3 else 131072, f[4] if len(f) > 4 else 4096, layers=4)
for layer in range(int(os.environ.get("LAYERS", "8"))):
    st.both(layer)   # the two launches ...
    st.fused(layer)  # ... and the one-launch layer (three kernels per iteration)
torch.cuda.synchronize()
print("ok", f)
