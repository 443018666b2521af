"""Experiment: per-CTA timeline of the fused score/select kernel (svb200_debug_times)."""
import os
import sys

import torch

from exp_common import DEV as dev, QuestState, ops
from sparse_vllm_b200 import _native

bsz, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '32,28,4').split(','))
st = QuestState(bsz, hq, hkv, int(os.environ.get('CTX', 131072)), int(os.environ.get('BUDGET', 4096)), layers=1, kv=False)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
buf = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
lib = _native.load()
for cfg in (sys.argv[1] if len(sys.argv) > 1 else "0:0:0").split(","):
    groups, minb, sdiv = (int(x) for x in cfg.split(":"))
    ops.set_tuning(0, groups); ops.set_tuning(2, minb); ops.set_tuning(3, sdiv); ops.set_tuning(4, int(os.environ.get('CHUNK', '0'))); ops.set_tuning(6, int(os.environ.get('QCOPY', '0')))
    buf.zero_()
    for it in range(4):
        flush.zero_()
        if it == 3:
            lib.svb200_debug_times(buf.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        st.score()
        e1.record(); torch.cuda.synchronize()
        if it >= 2:
            print("groups:minbatch:sharediv", cfg, "launch", it, round(e0.elapsed_time(e1) * 1e3, 1), "us")
    lib.svb200_debug_times(None)
    t = buf.view(148, 16).cpu()
    t0 = int(t[:, 0].min())
    rel = (t[:, :5] - t0).float() / 1000.0
    for name, col in (("first data", 2), ("consumers done", 3), ("producer done", 4)):
        c = rel[:, col]
        print(f"  {name:15s} min {c.min():7.2f} med {c.median():7.2f} max {c.max():7.2f}")
    if os.environ.get('CLAIMS'):
        lc = (t[bsz:, 15] - t0).float() / 1000.0
        dn = rel[bsz:, 3]
        drain = dn - lc
        order = dn.argsort()
        print("  last successful claim us: min", round(float(lc.min()), 1), "med", round(float(lc.median()), 1), "max", round(float(lc.max()), 1),
              "| done - last claim: min", round(float(drain.min()), 1), "med", round(float(drain.median()), 1), "max", round(float(drain.max()), 1))
        for i in list(order[:4]) + list(order[-6:]):
            print(f"    cta {int(i) + bsz:3d} sm {int(t[int(i) + bsz, 7]):3d} last claim {float(lc[i]):6.1f} done {float(dn[i]):6.1f} units {int(t[int(i) + bsz, 6])}")
    sel = t[:, 5].float() / 1000.0
    units = t[:, 6].float()
    w = t[:, 8:13].float().median(dim=0).values / 1000.0
    print("  median waits us: loader desc", round(float(w[0]), 1), "empty", round(float(w[1]), 1),
          "| consumer warp0 desc", round(float(w[3]), 1), "full", round(float(w[4]), 1))
    ph = t[:, 13:16].float() / 1000.0
    m = sel > 0
    if m.any():
        print("  select phases us (mean over selecting CTAs): load", round(float(ph[m, 0].mean()), 1), "search", round(float(ph[m, 1].mean()), 1), "compact+pack", round(float(ph[m, 2].mean()), 1))
    rate = units / (rel[:, 4] - rel[:, 2]).clamp(min=1.0)
    print("  select us: n>0", int((sel > 0).sum()), "mean", round(float(sel[sel > 0].mean()), 1), "max", round(float(sel.max()), 1),
          "| units/CTA min", int(units.min()), "med", int(units.median()), "max", int(units.max()), "sum", int(units.sum()),
          "| units/us while streaming: min", round(float(rate.min()), 2), "med", round(float(rate.median()), 2), "max", round(float(rate.max()), 2))
