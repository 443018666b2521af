"""Experiment: per-CTA timeline of the one-launch layer (svb200_debug_times, 32 slots per CTA)."""
import os

import torch

from exp_common import QuestState, ops
from sparse_vllm_b200 import _native

f = [int(x) for x in os.environ.get("SHAPE", "16,64,8").split(",")]
st = QuestState(f[0], f[1], f[2], f[3] if len(f) > 3 else 131072, f[4] if len(f) > 4 else 4096, layers=4)
lib = _native.load()
if os.environ.get("NOWAVES"):
    ops.set_tuning(12, int(os.environ["NOWAVES"]))
if os.environ.get("POLL"):
    ops.set_tuning(14, int(os.environ["POLL"]))
buf = torch.zeros(148 * 32, dtype=torch.int64, device="cuda:0")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for it in range(4):
    flush.zero_()
    if it == 3:
        lib.svb200_debug_times(buf.data_ptr())
    st.fused(it)
    torch.cuda.synchronize()
lib.svb200_debug_times(None)
t = buf.view(148, 32).cpu()
t = t[t[:, 0] > 0]
t0 = int(t[:, 0].min())
names = {0: "entry", 2: "score first data", 3: "score consumers done", 16: "phase 2 entry", 17: "attn set-up done",
         23: "first flag seen (producer 0)", 18: "attn first data", 20: "attn producer0 done", 19: "attn consumers done"}
print("CTAs", t.shape[0], "(us relative to the first CTA's entry; CTAs without a value are left out)")
for k, name in names.items():
    c = t[:, k]
    c = (c[c > 0] - t0).float() / 1000.0
    if c.numel():
        print(f"{name:30s} n {c.numel():3d} min {c.min():7.2f} med {c.median():7.2f} max {c.max():7.2f}")
ph = t[:, 13:16].float() / 1000.0
selecting = t[:, 5] > 0
if selecting.any():
    m = ph[selecting].mean(dim=0)
    print(f"select phases us (mean over selecting CTAs): load {m[0]:.1f} search {m[1]:.1f} compact+pack {m[2]:.1f}")
sel = t[:, 5].float() / 1000.0
if selecting.any():
    for i in torch.nonzero(selecting).flatten().tolist()[:16]:
        print(f"  selecting cta {i:3d}: score done {(int(t[i, 3]) - t0) / 1000.0:6.2f} select {float(sel[i]):5.1f} = load {float(ph[i, 0]):4.1f} search {float(ph[i, 1]):4.1f} "
              f"pack {float(ph[i, 2]):4.1f} | attn set-up done {(int(t[i, 17]) - t0) / 1000.0:6.2f} first data {(int(t[i, 18]) - t0) / 1000.0:6.2f}")
print("select us (selecting CTAs): mean", float(sel[sel > 0].mean()) if (sel > 0).any() else 0)
print("attn merge us: n", int((t[:, 21] > 0).sum()), "mean", float(t[:, 21][t[:, 21] > 0].float().mean() / 1000.0) if (t[:, 21] > 0).any() else 0,
      "| attn epilogue total us mean", float(t[:, 22].float().mean() / 1000.0))
