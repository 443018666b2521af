"""Experiment: per-CTA timeline of the fused score/select kernel (svb200_debug_times)."""
import sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, _native, configs
dev = "cuda:0"
import os
bsz, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '32,28,4').split(','))
ctx, budget = 131072, 4096
gen = torch.Generator(device=dev).manual_seed(1)
max_pages = ctx // 16
n_cols = max_pages - 1
base, prev = configs.quest_page_budget(budget, 16, max_pages)
num_pages = bsz * max_pages
bounds = torch.randn(num_pages, 2, hkv, 128, device=dev, generator=gen, dtype=torch.bfloat16)
pt = torch.randperm(num_pages, device=dev, generator=gen).to(torch.int32).view(bsz, max_pages)
q = torch.randn(bsz, hq, 128, device=dev, generator=gen).to(torch.bfloat16)
ctx_lens = torch.tensor([ctx - 3 * b for b in range(bsz)], dtype=torch.int32, device=dev)
req = torch.arange(bsz, dtype=torch.int32, device=dev)
scores = torch.empty(bsz, ops.score_row_stride(n_cols), dtype=torch.bfloat16, device=dev)
packed = torch.zeros(bsz, (prev + 1) * 16, dtype=torch.int32, device=dev)
lens = torch.zeros(bsz, dtype=torch.int32, device=dev)
tickets = ops.sched_scratch(bsz, dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
buf = torch.zeros(148 * 16, dtype=torch.int64, device=dev)
lib = _native.load()
import os
for cfg in (sys.argv[1] if len(sys.argv) > 1 else "0:0:0").split(","):
    groups, minb, sdiv = (int(x) for x in cfg.split(":"))
    ops.set_tuning(0, groups); ops.set_tuning(2, minb); ops.set_tuning(3, sdiv); ops.set_tuning(5, int(os.environ.get('SPLIT', '0'))); ops.set_tuning(6, int(os.environ.get('QCOPY', '0')))
    buf.zero_()
    for it in range(4):
        flush.zero_()
        if it == 3:
            lib.svb200_debug_times(buf.data_ptr())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.quest_score_select_pack(q, bounds, pt, req, ctx_lens, n_cols, hkv, prev, scores, packed, lens, tickets, None)
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
    print("  median waits us: producer0 desc", round(float(w[0]), 1), "empty", round(float(w[1]), 1), "ids", round(float(w[2]), 1),
          "| consumer warp0 desc", round(float(w[3]), 1), "full", round(float(w[4]), 1))
    ph = t[:, 13:16].float() / 1000.0
    m = sel > 0
    if m.any():
        print("  select phases us (mean over selecting CTAs): load", round(float(ph[m, 0].mean()), 1), "search", round(float(ph[m, 1].mean()), 1), "compact+pack", round(float(ph[m, 2].mean()), 1))
    rate = units / (rel[:, 4] - rel[:, 2]).clamp(min=1.0)
    print("  select us: n>0", int((sel > 0).sum()), "mean", round(float(sel[sel > 0].mean()), 1), "max", round(float(sel.max()), 1),
          "| units/CTA min", int(units.min()), "med", int(units.median()), "max", int(units.max()), "sum", int(units.sum()),
          "| units/us while streaming: min", round(float(rate.min()), 2), "med", round(float(rate.median()), 2), "max", round(float(rate.max()), 2))
