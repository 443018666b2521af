"""Experiment: per-CTA timeline of the attention kernel (svb200_debug_times)."""
import os
import sys
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops, _native
dev = "cuda:0"
B, hq, hkv = (int(x) for x in os.environ.get('SHAPE', '32,28,4').split(','))
sel_pages = 256
num_pages = 1 << 16
k = torch.randn(num_pages * 16, hkv, 128, device=dev, dtype=torch.bfloat16)
v = torch.randn(num_pages * 16, hkv, 128, device=dev, dtype=torch.bfloat16)
q = torch.randn(B, hq, 128, device=dev).to(torch.bfloat16)
out = torch.empty_like(q)
ws = ops.attn_workspace(B, hq, hkv, dev)
req = torch.arange(B, dtype=torch.int32, device=dev)
lens = torch.full((B,), sel_pages * 16, dtype=torch.int32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
pages = torch.stack([torch.randperm(num_pages, device=dev)[:sel_pages] for _ in range(B)])
table = (pages[:, :, None] * 16 + torch.arange(16, device=dev)[None, None]).reshape(B, -1).to(torch.int32).contiguous()
buf = torch.zeros(148 * 8, dtype=torch.int64, device=dev)
lib = _native.load()
for it in range(4):
    flush.zero_()
    if it == 3:
        lib.svb200_debug_times(buf.data_ptr())
    ops.sparse_decode_attn(q, k, v, table, req, lens, out=out, workspace=ws)
    torch.cuda.synchronize()
lib.svb200_debug_times(None)
t = buf.view(148, 8).cpu()
t = t[t[:, 0] > 0]  # CTAs of the launch (small launches use fewer than 148)
print("CTAs:", t.shape[0])
t0 = int(t[:, 0].min())
rel = (t[:, :5] - t0).float() / 1000.0
print("cols: entry, setup done, first data, consumers done, producer0 done (us, relative to first CTA entry)")
for name, col in (("entry", 0), ("setup", 1), ("first data", 2), ("consumers done", 3), ("producer done", 4)):
    c = rel[:, col]
    print(f"{name:15s} min {c.min():7.2f} med {c.median():7.2f} max {c.max():7.2f}")
cmb = t[:, 5].float() / 1000.0; epi = t[:, 6].float() / 1000.0
print("combine us: n>0", int((cmb > 0).sum()), "mean", float(cmb[cmb > 0].mean()) if (cmb > 0).any() else 0, "max", float(cmb.max()), "| epilogue total us: mean", float(epi.mean()), "max", float(epi.max()))
dur = rel[:, 3] - rel[:, 0]
print("corr(dur, epi):", float(torch.corrcoef(torch.stack([dur, epi]))[0, 1]))
order = dur.argsort()
print("fastest CTAs (cta, sm, dur):", [(int(i), int(t[i, 7]), round(float(dur[i]), 1)) for i in order[:6]])
print("slowest CTAs (cta, sm, dur):", [(int(i), int(t[i, 7]), round(float(dur[i]), 1)) for i in order[-6:]])
hist = torch.histc(dur, bins=10, min=float(dur.min()), max=float(dur.max()))
print("duration histogram", float(dur.min()), float(dur.max()), hist.tolist())
# second launch with other pages: are the same SMs slow (position) or is it the data (random pages)?
pages2 = torch.stack([torch.randperm(num_pages, device=dev)[:sel_pages] for _ in range(B)])
table2 = (pages2[:, :, None] * 16 + torch.arange(16, device=dev)[None, None]).reshape(B, -1).to(torch.int32).contiguous()
durs = {}
for name, tb in (("same", table), ("other", table2), ("same2", table)):
    flush.zero_()
    buf.zero_()
    lib.svb200_debug_times(buf.data_ptr())
    ops.sparse_decode_attn(q, k, v, tb, req, lens, out=out, workspace=ws)
    torch.cuda.synchronize()
    lib.svb200_debug_times(None)
    tt = buf.view(148, 8).cpu()
    tt = tt[tt[:, 0] > 0]
    d = (tt[:, 3] - tt[:, 0]).float() / 1000.0
    bysm = torch.zeros(148)
    bysm[tt[:, 7].long()] = d
    durs[name] = (d, bysm)
print("corr by CTA  first/same:", float(torch.corrcoef(torch.stack([dur, durs['same'][0]]))[0, 1]),
      " same/other:", float(torch.corrcoef(torch.stack([durs['same'][0], durs['other'][0]]))[0, 1]))
print("corr by SM   same/other:", float(torch.corrcoef(torch.stack([durs['same'][1], durs['other'][1]]))[0, 1]),
      " same/same2:", float(torch.corrcoef(torch.stack([durs['same'][1], durs['same2'][1]]))[0, 1]))
pd = (t[:, 4] - t[:, 0]).float() / 1000.0
if os.environ.get("VERBOSE"):
    print("producer0 done - by CTA index (first 148):", [round(float(x), 1) for x in pd.tolist()])
