"""Experiment: per-CTA timeline of the fused score/select kernel (svb200_debug_times)."""
import sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, _native, configs
dev = "cuda:0"
bsz, hq, hkv, ctx, budget = 32, 28, 4, 131072, 4096
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
tickets = torch.zeros(bsz, dtype=torch.int32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
buf = torch.zeros(148 * 8, dtype=torch.int64, device=dev)
lib = _native.load()
for it in range(4):
    flush.zero_()
    if it == 3:
        lib.svb200_debug_times(buf.data_ptr())
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    ops.quest_score_select_pack(q, bounds, pt, req, ctx_lens, n_cols, hkv, prev, scores, packed, lens, tickets, None)
    e1.record(); torch.cuda.synchronize()
    print("launch", it, e0.elapsed_time(e1) * 1e3, "us")
lib.svb200_debug_times(None)
t = buf.view(148, 8).cpu()
t0 = int(t[:, 0].min())
rel = (t[:, :5] - t0).float() / 1000.0
for name, col in (("entry", 0), ("setup", 1), ("first data", 2), ("consumers done", 3), ("producer done", 4)):
    c = rel[:, col]
    print(f"{name:15s} min {c.min():7.2f} med {c.median():7.2f} max {c.max():7.2f}")
sel = t[:, 5].float() / 1000.0
print("select time per CTA (us): n>0", int((sel > 0).sum()), "mean of those", float(sel[sel > 0].mean()), "max", float(sel.max()))
dur = rel[:, 3] - rel[:, 0]
print("duration histogram", float(dur.min()), float(dur.max()), torch.histc(dur, bins=10, min=float(dur.min()), max=float(dur.max())).tolist())
