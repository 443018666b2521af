"""Experiment: does the KV pool footprint (TLB reach) bound the sparse decode attention kernel?
Same launch (32 rows x 256 selected pages, 4 KV heads) over pools of different sizes."""
import sys, time
import torch
sys.path.insert(0, ".")
import svb200_import  # noqa
from sparse_vllm_b200 import ops

dev = "cuda:0"
B, hq, hkv, sel_pages = 32, 28, 4, 256
q = torch.randn(B, hq, 128, device=dev).to(torch.bfloat16)
out = torch.empty_like(q)
ws = ops.attn_workspace(B, hq, hkv, dev)
req = torch.arange(B, dtype=torch.int32, device=dev)
lens = torch.full((B,), sel_pages * 16, dtype=torch.int32, device=dev)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, n=20):
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2]


for log_pages in (14, 16, 18):
    num_pages = 1 << log_pages
    num_slots = num_pages * 16
    k = torch.randn(num_slots, hkv, 128, device=dev, dtype=torch.bfloat16)
    v = torch.randn(num_slots, hkv, 128, device=dev, dtype=torch.bfloat16)
    for mode in ("random", "sorted"):
        pages = torch.stack([torch.randperm(num_pages, device=dev)[:sel_pages] for _ in range(B)])
        if mode == "sorted":
            pages = pages.sort(dim=1).values
        table = (pages[:, :, None] * 16 + torch.arange(16, device=dev)[None, None]).reshape(B, -1).to(torch.int32).contiguous()
        ms = timeit(lambda: ops.sparse_decode_attn(q, k, v, table, req, lens, out=out, workspace=ws))
        nbytes = B * sel_pages * 16 * hkv * 256 * 2
        # torch gather of the same K pages (reads + writes the same bytes)
        kp = k.view(num_pages, -1)
        idx = pages.reshape(-1)
        ms_g = timeit(lambda: torch.index_select(kp, 0, idx))
        print(f"pool {num_slots * hkv * 256 / 2**20:8.0f} MiB/tensor {mode:6s}: attn {ms * 1e3:7.1f} us = {nbytes / ms / 1e6:7.0f} GB/s | "
              f"index_select K pages {ms_g * 1e3:7.1f} us = {nbytes / ms_g / 1e6:7.0f} GB/s (r+w)", flush=True)
    del k, v
