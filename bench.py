#!/usr/bin/env python
"""bench.py — sparse decode-attention tokens/s at 128K context (BASELINE.json metric).

A "step" is one decode step of the hot path over the whole batch: for each of the model's L layers,
Quest page scoring + top-k page selection + slot packing, then sparse paged GQA decode attention
(split-K, fused combine).  Every N runs BASELINE config 5 (Qwen3-32B shape, 64 q / 8 KV heads, bs 16,
128K ctx, budget 4096): at N>1 KV heads are sharded over the ranks with an NCCL all-gather of the
attention output per layer (strong scaling).  At N=1 config 3 (Qwen2.5-7B shape, bs 32, 128K ctx) is
measured as well and reported under "other_workloads".  See DESIGN.md "Measurement".

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--workload cfg2|cfg3|cfg5]
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path
from types import SimpleNamespace

ROOT = Path(__file__).resolve().parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

PAGE = 16
HEAD_DIM = 128
NOMINAL_GBS = 8000.0  # north_star: "B200's ~8 TB/s"
METRIC = "sparse_decode_attn_tokens_per_s_128k"
UNIT = "tok/s"

# BASELINE.json configs (shapes only; weights/data are synthetic).  budget = sink + decode_keep + recent.
WORKLOADS = {
    # tiny random-init Llama-style model (reference configs/debug/qwen3_tiny_random.json): CPU arms only
    "cfg1": dict(name="cfg1: tiny Llama-style model quest bs1 ctx4K budget512 (CPU reference path)", layers=2, hq=8, hkv=4,
                 bsz=1, ctx=4096, sink=64, decode_keep=384, recent=64),
    "cfg2": dict(name="cfg2: Llama-3.1-8B shape quest bs1 ctx32K budget2048", layers=32, hq=32, hkv=8, bsz=1,
                 ctx=32768, sink=64, decode_keep=1472, recent=512),
    "cfg3": dict(name="cfg3: Qwen2.5-7B shape quest bs32 ctx128K budget4096", layers=28, hq=28, hkv=4, bsz=32,
                 ctx=131072, sink=64, decode_keep=3520, recent=512),
    "cfg5": dict(name="cfg5: Qwen3-32B shape quest bs16 ctx128K budget4096 TP", layers=64, hq=64, hkv=8, bsz=16,
                 ctx=131072, sink=64, decode_keep=3520, recent=512),
    # h2o scored decode: the prompt is compacted to the budget by the prefill side, rows then cycle
    # budget+1 .. budget+interval tokens between two burst evictions
    "cfg4": dict(name="cfg4: Llama-3.1-8B shape h2o bs64 ctx32K budget2048 interval128", layers=32, hq=32, hkv=8,
                 bsz=64, ctx=32768, h2o_budget=2048, h2o_interval=128, sink=0, decode_keep=0, recent=0),
}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference", "incumbent"],
                    help="reference = the reference's CPU torch path on the host cores; incumbent = the reference's "
                         "own torch + Triton GPU path on this GPU (both from oracle/_ref, unmodified)")
    ap.add_argument("--workload", default="auto", choices=["auto", *WORKLOADS])
    ap.add_argument("--phys-layers", type=int, default=0, help="distinct KV/bounds pools cycled by the layers (0 = auto)")
    ap.add_argument("--skip-layers", type=int, default=0, help="quest_skip_layers (dense layers at the bottom)")
    ap.add_argument("--no-graph", action="store_true", help="time eager launches instead of a CUDA graph replay")
    ap.add_argument("--push-flags", action="store_true",
                    help="push epilogue with system fences + per-rank flags instead of the flag-in-data staging")
    ap.add_argument("--collective", default="push", choices=["push", "nccl"],
                    help="N>1: gather of the attention head shards: 'push' = peer stores in the attention epilogue "
                         "(svb200_sparse_decode_attn_tp), 'nccl' = all_gather_into_tensor after every layer")
    ap.add_argument("--one-launch", default="auto", choices=["auto", "on", "off"],
                    help="sparse layers as ONE launch (store+score+select+pack+attention, svb200_quest_decode_layer) "
                         "instead of two; auto = the provider's default")
    ap.add_argument("--pdl", type=int, default=-1, help="programmatic dependent launch: 1 on, 0 off, -1 library default")
    ap.add_argument("--tune", default="", help="experiment knobs 'which=value,...' (svb200_set_tuning)")
    ap.add_argument("--cpu-seconds", type=float, default=15.0, help="budget of the cpu_baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--e2e-chunks", type=int, default=2,
                    help="e2e arm: the step's layers run in this many chunks (a chunk's read-back overlaps the next chunk)")
    ap.add_argument("--verbose", action="store_true", help="phase timestamps on stderr")
    return ap.parse_args()


def pick_workload(args) -> dict:
    key = args.workload
    if key == "auto":
        # One workload at every N, so that the 1/2/4/8-GPU series is a strong-scaling curve: config 5
        # (8 KV heads, the largest configuration; it fits one B200).  Config 3 (4 KV heads, cannot be
        # sharded 8 ways) is measured next to it at N=1 and reported under "other_workloads".
        key = "cfg5"
    w = dict(WORKLOADS[key])
    w["key"] = key
    w["budget"] = w["sink"] + w["decode_keep"] + w["recent"]
    return w


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu_index = gpu_index
        self.rows: list[list[str]] = []
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.gpu_index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        except OSError:
            self.proc = None
            return
        def pump():
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
        self.thread = threading.Thread(target=pump, daemon=True)
        self.thread.start()

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        if self.thread is not None:
            self.thread.join(timeout=2)
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for r in self.rows:
            if len(r) < 7:
                continue
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except ValueError:
                continue
            for name, val in zip(names, r[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


# --------------------------------------------------------------------------------------- bytes
def geometry(w: dict, hkv_local: int, hq_local: int) -> dict:
    """Per (sequence, sparse layer, rank) algorithmic bytes — SURVEY.md §8(d) / DESIGN.md."""
    pages = (w["ctx"] + PAGE - 1) // PAGE
    base = max(3, w["budget"] // PAGE)
    prev = min(base - 1, pages - 1)
    t_sel = (prev + 1) * PAGE  # upper bound; the last page is partial by < 16 tokens
    select_bytes = 2 * (pages - 1) * hkv_local * HEAD_DIM * 2 + hq_local * HEAD_DIM * 2 + 4 * pages + 4 * (prev + 1) * PAGE
    attn_bytes = 2 * t_sel * hkv_local * HEAD_DIM * 2 + 2 * hq_local * HEAD_DIM * 2 + 4 * t_sel
    dense_bytes = 2 * w["ctx"] * hkv_local * HEAD_DIM * 2 + 2 * hq_local * HEAD_DIM * 2 + 4 * w["ctx"]
    return dict(pages=pages, prev=prev, t_sel=t_sel, select_bytes=select_bytes, attn_bytes=attn_bytes,
                dense_bytes=dense_bytes)


def ncu_traffic(workload_key: str, kernel: str):
    """DRAM bytes per launch of `kernel` from the committed ncu capture of this workload, or None."""
    try:
        return json.loads((ROOT / "profiles" / "r2_traffic.json").read_text())[workload_key][kernel]
    except Exception:
        return None


def measured_peak_gbs() -> tuple[float, str]:
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        try:
            return float(json.loads(p.read_text())["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# --------------------------------------------------------------------------------------- B200 arm
def build_state(w: dict, rank: int, world: int, phys_layers: int, skip_layers: int, headroom_steps: int = 0):
    """Cache manager with `phys` physical layers of KV + bounds filled THROUGH the provider's store path
    (save_rope_kv_if_needed -> store kernel -> on_kv_stored -> bounds kernel), rows at ~ctx tokens."""
    import torch
    import svb200_import  # noqa: F401
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.cache_manager.quest import QuestCacheManager
    from sparse_vllm_b200.utils_context import set_context

    dev = torch.device("cuda", torch.cuda.current_device())
    hkv, hq = w["hkv"] // world, w["hq"] // world
    cfg = configs.make_sparse_config(sparse_method="quest", sink_keep_tokens=w["sink"],
                                     decode_keep_tokens=w["decode_keep"], recent_keep_tokens=w["recent"],
                                     quest_skip_layers=skip_layers)
    pages_per_row = (w["ctx"] + PAGE - 1) // PAGE
    num_pages = w["bsz"] * pages_per_row
    layer_bytes = num_pages * PAGE * hkv * HEAD_DIM * 2 * 2 + num_pages * 2 * hkv * HEAD_DIM * 2
    if phys_layers <= 0:
        free, _ = torch.cuda.mem_get_info()
        phys_layers = int(max(2, min(w["layers"], 8, (free * 0.55) // layer_bytes)))
    mgr = QuestCacheManager(cfg, num_layers=phys_layers, num_kv_heads=hkv, num_pages=num_pages, max_rows=w["bsz"],
                            max_model_len=w["ctx"], device=dev)
    # physical pages in a seeded random order (non-contiguous page tables, like kernel_bench/test_decode.py)
    g = torch.Generator().manual_seed(20260609 + rank)
    mgr.free_pages_cpu_stack[:] = torch.randperm(num_pages, generator=g).numpy().astype("int32")
    # ... and bounds chunks in a shuffled order too (rows do not get consecutive chunks)
    mgr.free_chunks_cpu_stack[:] = torch.randperm(mgr.num_chunks, generator=g).numpy().astype("int32")
    headroom = max(256, (headroom_steps + 15) // 16 * 16)  # tokens the bench may append per row
    gen = torch.Generator(device=dev).manual_seed(1234 + rank)
    set_context(True)
    chunk = 1 << 17
    for b in range(w["bsz"]):
        n = w["ctx"] - headroom - (b * 7) % PAGE  # ragged: partial last pages
        done = 0
        while done < n:
            m = min(chunk, n - done)
            slots = mgr.allocate_prefill(b, m)
            for layer in range(phys_layers):
                k = torch.randn(m, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=gen)
                v = torch.randn(m, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=gen)
                mgr.save_rope_kv_if_needed(layer, k, v, slots)
            done += m
    set_context(False, is_long_text=True)
    mgr.kv_layer_map = lambda layer: layer % phys_layers  # L logical layers cycle over the physical pools
    torch.cuda.synchronize()
    return mgr, cfg, phys_layers, hq, hkv


class DecodeStep:
    """One decode step over all layers through the provider's public API (cache-manager hooks + backend)."""

    def __init__(self, w, mgr, phys_layers, hq, hkv, world, skip_layers, collective="push", push_flags=False):
        import torch
        from sparse_vllm_b200.attention_backend import B200AttentionBackend
        self.torch = torch
        self.w, self.mgr, self.phys, self.hq, self.hkv, self.world = w, mgr, phys_layers, hq, hkv, world
        self.skip = skip_layers
        dev = mgr.device
        L, B = w["layers"], w["bsz"]
        gen = torch.Generator(device=dev).manual_seed(77)
        self.q = torch.randn(L, B, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=gen)
        self.k_new = torch.randn(L, B, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=gen)
        self.v_new = torch.randn(L, B, hkv, HEAD_DIM, device=dev, dtype=torch.bfloat16, generator=gen)
        self.out = torch.zeros(L, B, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16)
        self.collective = collective if world > 1 else "none"
        self.comm = None
        self.gathered = None
        if world > 1:
            if self.collective == "push":
                from sparse_vllm_b200.parallel import PeerComm
                nb = B * hq * world * HEAD_DIM * 2
                self.comm = PeerComm(2 * nb, ll_bytes=0 if push_flags else None)  # two gathered buffers: a peer may be one layer ahead
                self.gathered = [self.comm.view(i * nb, (B, hq * world, HEAD_DIM), torch.bfloat16) for i in range(2)]
            else:
                self.gathered = [torch.zeros(world, B, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16) for _ in range(2)]
        self.backend = B200AttentionBackend(provider=mgr._get_provider(hq))
        self.input_ids = torch.zeros(B, dtype=torch.int64, device=dev)
        self.positions = torch.zeros(B, dtype=torch.int64, device=dev)
        self.slot_mapping = torch.zeros(B, dtype=torch.int32, device=dev)
        self.context_lens = torch.zeros(B, dtype=torch.int32, device=dev)
        self.req_indices = torch.zeros(B, dtype=torch.int32, device=dev)
        self.seqs = [SimpleNamespace(seq_id=b, decode_input_token=1, decode_input_position=0) for b in range(B)]
        mgr.set_decode_static_max_context_len(w["ctx"])
        # pinned host staging for the end-to-end arm
        self.h_q = torch.empty(self.q.shape, dtype=torch.bfloat16).pin_memory().copy_(self.q.cpu())
        self.h_k = torch.empty(self.k_new.shape, dtype=torch.bfloat16).pin_memory().copy_(self.k_new.cpu())
        self.h_v = torch.empty(self.v_new.shape, dtype=torch.bfloat16).pin_memory().copy_(self.v_new.cpu())
        self.h_out = torch.empty(self.out.shape, dtype=torch.bfloat16).pin_memory()
        self.launches_per_step = 0
        self.timers = None  # optional per-kernel CUDA-event collection
        self.force_two_launches = False  # per-kernel timing: the two launches of a sparse layer, each alone
        self.no_comm = False         # per-kernel timing: the launches without their collective ...
        self.only_collective = False  # ... and (NCCL arm) the collective without the launches

    def one_launch_active(self) -> bool:
        """Does the manager run a sparse layer of this workload as one launch (its "auto" rule included)?"""
        return self.one_launch_active_with(getattr(self.mgr, "fuse_decode_layer", False))

    def one_launch_active_with(self, f) -> bool:
        mgr = self.mgr
        if not f or self.hq // self.hkv > 8:
            return False
        if f == "auto":
            bounds_bytes = self.w["bsz"] * mgr.max_pages_per_row * 2 * mgr.num_kv_heads * 128 * 2
            return bounds_bytes >= 148 * (3 << 19)
        return True

    def prepare(self):
        """Host side of a decode step: allocate one slot per row, fill the static metadata buffers."""
        self.mgr.prepare_decode_static(self.seqs, self.input_ids, self.positions, self.slot_mapping,
                                       self.context_lens, self.req_indices)

    def layers(self, lo: int = 0, hi: int | None = None):
        """Device side (layers [lo, hi), default all): per layer hand the new K/V row to the cache manager (stored, bounds refresh included,
        by the layer's scoring launch), build the Quest view, run decode attention; under TP all-gather the
        head shards."""
        from sparse_vllm_b200.attention_backend import decode_attention_layer
        from sparse_vllm_b200.cache_manager.base import SparseSelection
        torch, mgr = self.torch, self.mgr
        st = mgr.layer_batch_state
        n = 0
        hi = self.w["layers"] if hi is None else hi
        for layer in range(lo, hi):
            mgr.save_rope_kv_if_needed(layer, self.k_new[layer], self.v_new[layer], st.slot_mapping)
            sel = SparseSelection(kind="full", req_indices=st.req_indices, context_lens=st.context_lens,
                                  max_context_len=st.max_context_len)
            push = self.collective == "push" and not self.no_comm
            if self.only_collective:
                import torch.distributed as dist
                dist.all_gather_into_tensor(self.gathered[layer & 1].view(-1), self.out[layer].view(-1))
                continue
            decode_attention_layer(mgr, self.backend, layer, self.q[layer], sel, num_heads=self.hq,
                                   num_kv_heads=self.hkv, out=self.out[layer], tp_comm=self.comm if push else None,
                                   tp_gathered=self.gathered[layer & 1] if push else None)
            # dense layer: store + attention; sparse layer: (store fused into) score/select/pack + attention
            # dense layer: store + attention; sparse layer: (store fused into) score/select/pack + attention, or ONE launch
            one = layer >= self.skip and not self.force_two_launches and self.one_launch_active()
            n += 1 if one else (2 if (layer < self.skip or getattr(mgr, "fuse_decode_store", False)) else 3)
            if self.collective == "nccl" and not self.no_comm:
                import torch.distributed as dist
                dist.all_gather_into_tensor(self.gathered[layer & 1].view(-1), self.out[layer].view(-1))
        if lo == 0 and hi == self.w["layers"] and not (self.force_two_launches or self.no_comm or self.only_collective):
            self.launches_per_step = n  # the timed step's own count, not a per-kernel timing pass's


def verify_gather(step) -> None:
    """N>1, before anything is timed: after one eager step the gathered attention output of the last layer must
    hold every rank's head shard at its place — checked against an NCCL all-gather of the ranks' local outputs
    (bit for bit), on every rank."""
    import torch
    import torch.distributed as dist
    L, world = step.w["layers"], step.world
    layer = L - 1
    local = step.out[layer].contiguous()
    ref = torch.empty(world, *local.shape, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(ref.view(-1), local.view(-1))
    got = step.gathered[layer & 1]
    if step.collective == "push":  # [B, world*Hq, D] with rank r's heads at r*Hq
        want = ref.permute(1, 0, 2, 3).reshape(got.shape)
    else:
        want = ref
    if not torch.equal(got, want):
        bad = (got != want).nonzero()[:4].tolist()
        raise SystemExit(f"rank {dist.get_rank()}: gathered attention output differs from the ranks' shards at {bad}")
    # every rank also holds the same gathered tensor (cross-check through a MAX/MIN all-reduce of a checksum)
    chk = got.view(torch.int16).to(torch.int64).sum().reshape(1)
    lo, hi = chk.clone(), chk.clone()
    dist.all_reduce(lo, op=dist.ReduceOp.MIN)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX)
    if int(lo) != int(hi):
        raise SystemExit("ranks disagree on the gathered attention output")


def world_of(step) -> int:
    return int(step.world)


def graph_kernel_times(step, stream, lo: int, hi: int, reps: int = 5) -> dict:
    """Average duration of the score+select launch and of the attention launch of layers [lo, hi): each is
    captured ALONE in a CUDA graph (the other one replaced by a no-op) and replayed, so the figure is the
    kernel's own time back to back with itself — no Python launch cadence in it (the eager event pairs of
    round 1 measured the host at small per-rank work)."""
    import torch
    prov = step.mgr._get_provider(step.hq)
    real_select, real_decode = prov.quest_select, prov.decode
    out = {}
    # N > 1: the two launches are timed without the collective; then the collective's share of a layer: the attention
    # launch WITH its push epilogue (every rank replays the same graph, so the hand-shake is the real one), or the
    # NCCL all-gather alone
    keys = ["quest_score_select_pack", "sparse_decode_attn"]
    if step.collective == "push":
        keys.append("sparse_decode_attn+push")
    elif step.collective == "nccl":
        keys.append("nccl_all_gather")
    fuse_was = getattr(step.mgr, "fuse_decode_layer", False)
    try:
        for key in keys + (["quest_decode_layer"] if step.hq // step.hkv <= 8 and world_of(step) == 1 else []):
            # the two launches of a sparse layer, each alone; then (when it is the path in use) the one-launch layer
            step.mgr.fuse_decode_layer = key == "quest_decode_layer"  # (True: forced, whatever the "auto" rule says)
            step.force_two_launches = not step.mgr.fuse_decode_layer
            if key == "quest_decode_layer":
                prov.quest_select, prov.decode = real_select, real_decode
                step.no_comm, step.only_collective = True, False
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    step.layers(lo, hi)
                g.replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream)
                for _ in range(reps):
                    g.replay()
                e1.record(stream)
                stream.synchronize()
                n = reps * (hi - lo)
                out[key] = {"launches": n, "avg_ms": e0.elapsed_time(e1) / n, "how": "graph replay of this launch alone"}
                g.reset()
                continue
            prov.quest_select = real_select if key == "quest_score_select_pack" else (lambda *a, **k: None)
            prov.decode = real_decode if key.startswith("sparse_decode_attn") else (lambda q, *a, out=None, **k: out)
            step.no_comm = key in ("quest_score_select_pack", "sparse_decode_attn")
            step.only_collective = key == "nccl_all_gather"
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g, stream=stream):
                step.layers(lo, hi)
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(stream)
            for _ in range(reps):
                g.replay()
            e1.record(stream)
            stream.synchronize()
            n = reps * (hi - lo)
            out[key] = {"launches": n, "avg_ms": e0.elapsed_time(e1) / n, "how": "graph replay of this launch alone"}
            g.reset()
    finally:
        prov.quest_select, prov.decode = real_select, real_decode
        step.no_comm = step.only_collective = False
        step.mgr.fuse_decode_layer = fuse_was
        step.force_two_launches = False
    return out


class KernelTimers:
    """CUDA-event pairs around the provider's two hot launches (recorded on the launching stream)."""

    def __init__(self, provider):
        import torch
        self.torch = torch
        self.pairs = {"quest_score_select_pack": [], "sparse_decode_attn": [], "quest_decode_layer": []}
        self.enabled = False
        self._select, self._decode = provider.quest_select, provider.decode
        provider.quest_decode_layer = self._wrap(provider.quest_decode_layer, "quest_decode_layer")
        provider.quest_select = self._wrap(self._select, "quest_score_select_pack")
        provider.decode = self._wrap(self._decode, "sparse_decode_attn")

    def _wrap(self, fn, key):
        def inner(*a, **k):
            if not self.enabled:
                return fn(*a, **k)
            e0 = self.torch.cuda.Event(enable_timing=True)
            e1 = self.torch.cuda.Event(enable_timing=True)
            e0.record()
            r = fn(*a, **k)
            e1.record()
            self.pairs[key].append((e0, e1))
            return r
        return inner

    def summary(self) -> dict:
        out = {}
        self.torch.cuda.synchronize()  # the event pairs were recorded on a side stream
        for key, pairs in self.pairs.items():
            ms = [a.elapsed_time(b) for a, b in pairs]
            if ms:
                out[key] = {"launches": len(ms), "avg_ms": sum(ms) / len(ms), "total_ms": sum(ms),
                            "how": "eager launches, CUDA event pairs"}
        return out


def run_b200(args) -> dict:
    import torch
    import torch.distributed as dist

    t_start = time.perf_counter()
    if args.verbose:
        import faulthandler
        faulthandler.dump_traceback_later(75, exit=False, file=sys.stderr)

    def note(msg):
        if args.verbose:
            print(f"[bench rank {os.environ.get('RANK', '0')} +{time.perf_counter() - t_start:6.1f}s] {msg}",
                  file=sys.stderr, flush=True)

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        if world == 1 and args.gpus > 1:
            raise SystemExit("--gpus N>1 must be launched with torch.distributed.run (one rank per GPU).")
        raise SystemExit(f"--gpus {args.gpus} but WORLD_SIZE={world}")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    note("process group ready")
    w = pick_workload(args)
    if w["hkv"] % world:
        raise SystemExit(f"{w['key']}: {w['hkv']} KV heads cannot be sharded over {world} ranks")

    mgr, cfg, phys, hq, hkv = build_state(w, rank, world, args.phys_layers, args.skip_layers,
                                           headroom_steps=2 * (args.steps + max(args.warmup, 3)) + 16)
    note(f"state built ({phys} physical layers)")
    from sparse_vllm_b200 import ops as _ops
    if args.pdl >= 0:
        _ops.set_tuning(7, 1 if args.pdl else 2)
    for item in filter(None, args.tune.split(",")):
        which, value = item.split("=")
        _ops.set_tuning(int(which), int(value))
    if args.one_launch != "auto":
        mgr.fuse_decode_layer = args.one_launch == "on"
    step = DecodeStep(w, mgr, phys, hq, hkv, world, args.skip_layers, collective=args.collective, push_flags=args.push_flags)
    geo = geometry(w, hkv, hq)
    steps, warmup = args.steps, max(args.warmup, 3)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- (1) device-resident throughput: inputs already in HBM, CUDA-graph replay of one step
    step.prepare()
    barrier()  # ranks reach their first collective launch together
    stream = torch.cuda.Stream()
    graph = None
    with torch.cuda.stream(stream):
        for _ in range(2):
            step.layers()
        stream.synchronize()
        if world > 1:
            verify_gather(step)
        note("eager steps done")
        if not args.no_graph:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph, stream=stream):
                step.layers()
        run = graph.replay if graph is not None else step.layers
        note("graph captured" if graph is not None else "no graph")
        for _ in range(warmup):
            run()
        barrier()
        note("warm-up done")
        sampler = ClockSampler(local_rank)
        sampler.start()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            run()
        e1.record(stream)
        barrier()
        dev_ms = e0.elapsed_time(e1)
        note("timed region done")

        # ---- (2) per-kernel durations, live: each hot launch alone in a graph (eager event pairs with --no-graph)
        skip = args.skip_layers
        kdense = None
        if graph is not None:
            ksum = graph_kernel_times(step, stream, skip, w["layers"])
            if skip > 0:
                kdense = graph_kernel_times(step, stream, 0, skip)["sparse_decode_attn"]
        else:
            timers = KernelTimers(mgr._get_provider(hq))
            step.layers()
            timers.enabled = True
            for _ in range(min(steps, 5)):
                step.layers(skip)
            fuse_was = getattr(mgr, "fuse_decode_layer", False)
            mgr.fuse_decode_layer = False  # ... and the two launches of a sparse layer, whatever path the step uses
            for _ in range(min(steps, 5) if step.one_launch_active_with(fuse_was) else 0):
                step.layers(skip)
            mgr.fuse_decode_layer = fuse_was
            timers.enabled = False
            ksum = timers.summary()
        barrier()
        note("per-kernel timing done")

        # ---- (3) end to end through the public API with HOST buffers: every step copies its q / new K,V
        # rows from pinned host memory and reads the attention output back.  The layers run in chunks so that
        # a chunk's read-back (third stream) overlaps the next chunk's compute; input copies run on a copy stream.
        L = w["layers"]
        n_chunks = args.e2e_chunks if L % max(args.e2e_chunks, 1) == 0 and args.e2e_chunks >= 1 else 1
        per = L // n_chunks
        copy_s, back_s = torch.cuda.Stream(), torch.cuda.Stream()
        # Two sets of device input buffers: step i + 1's inputs are copied in while step i computes (a serving loop
        # double-buffers its inputs the same way); the chunk graphs are captured once per set.
        sets = [(step.q, step.k_new, step.v_new),
                (torch.empty_like(step.q), torch.empty_like(step.k_new), torch.empty_like(step.v_new))]
        chunk_graphs = [[], []]
        if graph is not None:
            for si, (q_, k_, v_) in enumerate(sets):
                step.q, step.k_new, step.v_new = q_, k_, v_
                for c in range(n_chunks):
                    g = torch.cuda.CUDAGraph()
                    with torch.cuda.graph(g, stream=stream):
                        step.layers(c * per, (c + 1) * per)
                    chunk_graphs[si].append(g)
        in_ready = [[torch.cuda.Event() for _ in range(n_chunks)] for _ in sets]
        out_ready = [torch.cuda.Event() for _ in range(n_chunks)]
        d2h_done = [torch.cuda.Event() for _ in range(n_chunks)]
        set_free = [torch.cuda.Event() for _ in sets]
        e2e_count = [0]
        host_t = [0.0, 0.0, 0.0]  # host seconds in prepare / input copies / launches (verbose note)

        def e2e_step():
            i = e2e_count[0]
            e2e_count[0] += 1
            si = i & 1
            step.q, step.k_new, step.v_new = sets[si]
            tp = time.perf_counter()
            step.prepare()
            host_t[0] += time.perf_counter() - tp
            tp = time.perf_counter()
            if i >= 2:
                copy_s.wait_event(set_free[si])  # step i - 2 has read this set of q / k / v
            with torch.cuda.stream(copy_s):
                for c in range(n_chunks):
                    sl = slice(c * per, (c + 1) * per)
                    step.q[sl].copy_(step.h_q[sl], non_blocking=True)
                    step.k_new[sl].copy_(step.h_k[sl], non_blocking=True)
                    step.v_new[sl].copy_(step.h_v[sl], non_blocking=True)
                    in_ready[si][c].record(copy_s)
            host_t[1] += time.perf_counter() - tp
            for c in range(n_chunks):
                sl = slice(c * per, (c + 1) * per)
                stream.wait_event(in_ready[si][c])
                if i >= 1:
                    stream.wait_event(d2h_done[c])  # the previous step's output of this chunk is on the host
                tp = time.perf_counter()
                if chunk_graphs[si]:
                    chunk_graphs[si][c].replay()
                else:
                    step.layers(c * per, (c + 1) * per)
                host_t[2] += time.perf_counter() - tp
                out_ready[c].record(stream)
                back_s.wait_event(out_ready[c])
                with torch.cuda.stream(back_s):
                    step.h_out[sl].copy_(step.out[sl], non_blocking=True)
                    d2h_done[c].record(back_s)
            set_free[si].record(stream)

        def e2e_drain():
            stream.wait_stream(back_s)  # the last step's result is on the host
        for _ in range(warmup):
            e2e_step()
        e2e_drain()
        barrier()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        host_t[:] = [0.0, 0.0, 0.0]
        t_host = time.perf_counter()
        for _ in range(steps):
            e2e_step()
        t_host = time.perf_counter() - t_host  # host time to enqueue the steps (no sync inside: the host runs ahead)
        e2e_drain()
        f1.record(stream)
        barrier()
        e2e_ms = f0.elapsed_time(f1)
        note(f"e2e done: host enqueue {t_host / steps * 1e3:.3f} ms/step (prepare {host_t[0] / steps * 1e3:.3f}, input copies {host_t[1] / steps * 1e3:.3f}, launches {host_t[2] / steps * 1e3:.3f}), device {e2e_ms / steps:.3f} ms/step")
        clocks = sampler.stop()
        note("clock sampler stopped")

    times = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    dev_ms, e2e_ms = (float(x) for x in times.tolist())
    ms_per_step = dev_ms / steps
    value = w["bsz"] / (ms_per_step * 1e-3)
    e2e_value = w["bsz"] / (e2e_ms / steps * 1e-3)
    h2d = step.h_q.numel() * 2 + step.h_k.numel() * 2 + step.h_v.numel() * 2 + 3 * 4 * w["bsz"]
    d2h = step.h_out.numel() * 2

    peak, peak_src = measured_peak_gbs()
    sel_ms = ksum["quest_score_select_pack"]["avg_ms"]
    attn_ms = ksum["sparse_decode_attn"]["avg_ms"]
    sel_bytes = geo["select_bytes"] * w["bsz"]
    attn_bytes = geo["attn_bytes"] * w["bsz"]
    sparse_layers = w["layers"] - args.skip_layers
    step_bytes = (sel_bytes + attn_bytes) * sparse_layers + geo["dense_bytes"] * w["bsz"] * args.skip_layers
    roof = {"bound": "hbm", "kernel": "quest_page_score_kernel<fused select+pack>", "achieved": sel_bytes / (sel_ms * 1e-3) / 1e9,
            "peak": peak, "unit": "GB/s", "peak_source": peak_src,
            "traffic": ncu_traffic(w["key"], "quest_page_score_kernel") if world == 1 else None,
            "bytes_per_launch": sel_bytes, "avg_launch_ms": sel_ms}
    roof["frac"] = roof["achieved"] / peak
    roof["how"] = ksum["quest_score_select_pack"].get("how", "eager launches, CUDA event pairs")
    roof_attn = {"bound": "hbm", "kernel": "sparse_decode_attn_kernel", "achieved": attn_bytes / (attn_ms * 1e-3) / 1e9,
                 "peak": peak, "unit": "GB/s", "bytes_per_launch": attn_bytes, "avg_launch_ms": attn_ms,
                 "traffic": ncu_traffic(w["key"], "sparse_decode_attn_kernel") if world == 1 else None}
    roof_attn["frac"] = roof_attn["achieved"] / peak
    roof_attn["how"] = ksum["sparse_decode_attn"].get("how", "eager launches, CUDA event pairs") + (
        " (without the collective: see `collective`)" if world > 1 else "")
    roof_step = {"achieved": step_bytes / (ms_per_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                 "bytes_per_step": step_bytes}
    roof_step["frac"] = roof_step["achieved"] / peak
    one_launch = None
    if "quest_decode_layer" in ksum:
        t = ksum["quest_decode_layer"]["avg_ms"]
        one_launch = {"kernel": "quest_page_score_kernel<one-launch layer: store+score+select+pack+attention>",
                      "avg_launch_ms": t, "bytes_per_launch": sel_bytes + attn_bytes,
                      "achieved": (sel_bytes + attn_bytes) / (t * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                      "frac": (sel_bytes + attn_bytes) / (t * 1e-3) / 1e9 / peak,
                      "two_launches_ms": sel_ms + attn_ms, "how": "graph replay of this launch alone"}
    collective = None
    if world > 1 and "sparse_decode_attn+push" in ksum:
        with_push = ksum["sparse_decode_attn+push"]["avg_ms"]
        collective = {"kind": "push epilogue of the attention launch (peer stores over NVLink + flag hand-shake)",
                      "us_per_layer": (with_push - attn_ms) * 1e3, "attention_with_push_ms": with_push}
    elif world > 1 and "nccl_all_gather" in ksum:
        collective = {"kind": "ncclAllGather of the head shards, one per layer",
                      "us_per_layer": ksum["nccl_all_gather"]["avg_ms"] * 1e3}
    if collective is not None:
        layer_us = ms_per_step * 1e3 / w["layers"]
        collective["layer_us"] = layer_us
        collective["other_us_per_layer"] = layer_us - (sel_ms + attn_ms) * 1e3 - collective["us_per_layer"]
        collective["other_is"] = "launch gaps of the graph + waiting for the slowest rank of the layer"
    for r_ in (roof, roof_attn, roof_step):  # north_star quotes ~8 TB/s nominal: report against that too
        r_["frac_of_nominal_8tbs"] = r_["achieved"] / NOMINAL_GBS
    roof_dense = None
    if kdense is not None:
        dense_bytes = geo["dense_bytes"] * w["bsz"]
        roof_dense = {"bound": "hbm", "kernel": "sparse_decode_attn_kernel (dense skip layer, identity slot table)",
                      "achieved": dense_bytes / (kdense["avg_ms"] * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                      "bytes_per_launch": dense_bytes, "avg_launch_ms": kdense["avg_ms"]}
        roof_dense["frac"] = roof_dense["achieved"] / peak
        roof_dense["frac_of_nominal_8tbs"] = roof_dense["achieved"] / NOMINAL_GBS
    def shutdown():
        """Release the captured graph BEFORE the NCCL communicator it references (destroying the process
        group under a live graph hangs), and never let clean-up keep the job alive."""
        nonlocal graph
        if world == 1:
            return
        watchdog = threading.Timer(30.0, os._exit, (0,))
        watchdog.daemon = True
        watchdog.start()
        if graph is not None:
            graph.reset()
            graph = None
        for gs in chunk_graphs:
            for g in gs:
                g.reset()
            gs.clear()
        torch.cuda.synchronize()
        dist.barrier()
        if step.comm is not None:
            step.comm.close()
        dist.destroy_process_group()
        watchdog.cancel()

    if rank != 0:
        shutdown()
        return {}
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": w["name"], "layers": w["layers"], "q_heads": w["hq"], "kv_heads": w["hkv"],
                   "batch": w["bsz"], "ctx": w["ctx"], "page": PAGE, "token_budget": w["budget"],
                   "quest_skip_layers": args.skip_layers,
                   "parallelism": f"tp{world} (KV heads sharded"
                   + (f"; head shards gathered by {('peer stores in the attention epilogue (NVLink), ' + ('fences + flags' if args.push_flags else 'flag-in-data words')) if args.collective == 'push' else 'NCCL all_gather_into_tensor'})" if world > 1 else ")"),
                   "pdl": {1: "on", 0: "off"}.get(args.pdl, "library default"),
                   "physical_kv_pools": phys, "l2": "inputs larger than L2 (>= 0.8 GB touched per layer; layers cycle "
                   f"{phys} distinct KV+bounds pools)",
                   "timed": ("cuda graph replay of one decode step (slots and lengths frozen at the step prepared before "
                             "capture; only the e2e arm advances the context)") if graph is not None else "eager"},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": e2e_ms / steps,
                "how": f"{n_chunks} chunks of {per} layers; a chunk's H2D copy (two sets of device input buffers: the next "
                       "step's inputs arrive while this one computes) and the previous chunk's D2H read-back run on side "
                       "streams next to compute, all inside the timed region, which ends when the last result is on the host"},
        "gpu_launches": step.launches_per_step * steps,
        "roofline": roof, "roofline_attention": roof_attn, "roofline_step": roof_step,
        "launches_per_sparse_layer": 1 if step.one_launch_active() else 2,
        "clocks": clocks,
    }
    if roof_dense is not None:
        line["roofline_dense_layer"] = roof_dense
    if collective is not None:
        line["collective"] = collective
    if one_launch is not None:
        one_launch.update({"bound": "hbm", "peak_source": peak_src, "frac_of_nominal_8tbs": one_launch["achieved"] / NOMINAL_GBS,
                           "traffic": ncu_traffic(w["key"], "quest_decode_layer_kernel") if world == 1 else None})
        if step.one_launch_active():
            # the kernel that runs in the timed step is the one-launch layer: it is the dominant kernel of the line;
            # its two halves, each timed alone as separate launches, explain it
            line["roofline"], line["roofline_score_select_pack"] = one_launch, roof
        else:
            line["roofline_one_launch_layer"] = one_launch
    if world > 1:
        print(json.dumps(line), flush=True)  # before clean-up: the line must not depend on it
        line["_printed"] = True
        shutdown()
    return line


# --------------------------------------------------------------------------------------- h2o arm
def run_h2o(args) -> dict:
    """cfg4: H2O scored decode through the cache-manager hooks, steady state over whole eviction
    intervals (every step: per-layer slot claim, KV store, scored attention, normalise + accumulate;
    one burst eviction per interval).  A "step" is one decode step of the batch over all layers."""
    import torch
    from types import SimpleNamespace
    import svb200_import  # noqa: F401
    from sparse_vllm_b200 import configs
    from sparse_vllm_b200.attention_backend import B200AttentionBackend, decode_attention_layer
    from sparse_vllm_b200.cache_manager.base import CacheManager
    from sparse_vllm_b200.utils_context import reset_context, set_context

    if args.gpus != 1:
        raise SystemExit("cfg4 is a single-GPU workload")
    w = pick_workload(args)
    L, hq, hkv, bsz, budget, interval = w["layers"], w["hq"], w["hkv"], w["bsz"], w["h2o_budget"], w["h2o_interval"]
    width = budget + interval
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    reset_context()
    set_context(False)
    chunk, window = 4096, 128
    cfg = configs.make_sparse_config(sparse_method="h2o", h2o_decode_budget=budget, h2o_prefill_budget=budget,
                                     h2o_decode_eviction_interval=interval, h2o_recent_ratio=0.5,
                                     h2o_prefill_score_window=window)
    mgr = CacheManager.create(cfg, num_layers=L, num_kv_heads=hkv, num_slots=bsz * (budget + chunk) + 64, max_rows=bsz,
                              device=dev, max_prefill_chunk=chunk)
    gen = torch.Generator(device=dev).manual_seed(11)
    # ---- the 32K prompt of every row goes through the provider's prefill side (h2o.py:665-894, 1352-1496): chunks of
    # 4096 tokens are appended to the compressed row of every layer (KV store), the last 128 queries of the chunk score
    # the row (svb200_h2o_prefill_score, cumulative scores accumulate), rows over the budget are compacted after every
    # chunk (svb200_h2o_select_compact).  Activations are synthetic per layer, so the dense prefill attention
    # (svb200_prefill_attn) has no consumer here and is not run.
    set_context(True)
    ids_all = list(range(bsz))
    n_chunks = w["ctx"] // chunk
    k_src = torch.randn(bsz * chunk, hkv, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    v_src = torch.randn(bsz * chunk, hkv, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    q_src = torch.randn(bsz * chunk, hq, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    t_pre0 = time.perf_counter()
    for c in range(n_chunks):
        slots = torch.stack([mgr.allocate_prefill_chunk(b, chunk) for b in ids_all], dim=1).reshape(L, bsz * chunk).contiguous()
        for layer in range(L):
            mgr.save_rope_kv_if_needed(layer, k_src, v_src, slots[layer])
            mgr.collect_prefill_attention_score(layer, q_src, ids_all, [chunk] * bsz)
        mgr.evict_after_prefill(ids_all, final=c == n_chunks - 1)
        k_src.mul_(-1.0)  # cheap variation between chunks
    torch.cuda.synchronize()
    prefill_s = time.perf_counter() - t_pre0
    del k_src, v_src, q_src, slots
    set_context(False)
    backend = B200AttentionBackend(provider=mgr._get_provider(hq))
    seqs = [SimpleNamespace(seq_id=b) for b in range(bsz)]
    ids = torch.zeros(bsz, dtype=torch.int64, device=dev)
    pos = torch.zeros(bsz, dtype=torch.int64, device=dev)
    slot_mapping = torch.zeros(bsz, dtype=torch.int32, device=dev)
    ctx = torch.zeros(bsz, dtype=torch.int32, device=dev)
    req = torch.zeros(bsz, dtype=torch.int32, device=dev)
    q = torch.randn(L, bsz, hq, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    k_new = torch.randn(L, bsz, hkv, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    v_new = torch.randn(L, bsz, hkv, HEAD_DIM, generator=gen, device=dev, dtype=torch.bfloat16)
    out = torch.empty(L, bsz, hq, HEAD_DIM, device=dev, dtype=torch.bfloat16)
    h_q, h_k, h_v = (torch.empty(t.shape, dtype=t.dtype).pin_memory().copy_(t) for t in (q, k_new, v_new))
    h_out = torch.empty(out.shape, dtype=out.dtype).pin_memory()

    # two sets of device buffers: in the end-to-end arm step i + 1's inputs are copied in (and step i's output is read
    # back) on side streams while the layers of the other set run
    sets = [(q, k_new, v_new, out), tuple(torch.empty_like(t) for t in (q, k_new, v_new, out))]

    def layers(si: int = 0):
        q_, k_, v_, o_ = sets[si]
        for layer in range(L):
            mgr.save_rope_kv_if_needed(layer, k_[layer], v_[layer])
            decode_attention_layer(mgr, backend, layer, q_[layer], mgr.decode_selection(layer), num_heads=hq,
                                   num_kv_heads=hkv, out=o_[layer])

    stream = torch.cuda.Stream()
    evictions = 0
    with torch.cuda.stream(stream):
        mgr.prepare_decode_static(seqs, ids, pos, slot_mapping, ctx, req)
        layers()
        evictions += bool(mgr.evict_after_decode())
        stream.synchronize()
        graph = None
        graphs = []
        if not args.no_graph:
            for si in range(2):
                mgr.prepare_decode_static(seqs, ids, pos, slot_mapping, ctx, req)
                g = torch.cuda.CUDAGraph()
                with torch.cuda.graph(g, stream=stream):
                    layers(si)
                g.replay()  # capture does not execute: this is the step the prepare above belongs to
                evictions += bool(mgr.evict_after_decode())
                graphs.append(g)
            graph = graphs[0]
        run_layers = (lambda si=0: graphs[si].replay()) if graphs else layers
        copy_s, back_s = torch.cuda.Stream(), torch.cuda.Stream()
        in_ready = [torch.cuda.Event() for _ in sets]
        layers_done = [torch.cuda.Event() for _ in sets]
        d2h_done = [torch.cuda.Event() for _ in sets]
        e2e_count = [0]

        def step(e2e: bool):
            nonlocal evictions
            mgr.prepare_decode_static(seqs, ids, pos, slot_mapping, ctx, req)
            if not e2e:
                run_layers(0)
                evictions += bool(mgr.evict_after_decode())
                return
            i = e2e_count[0]
            e2e_count[0] += 1
            si = i & 1
            q_, k_, v_, o_ = sets[si]
            if i >= 2:
                copy_s.wait_event(layers_done[si])  # step i - 2 has read this set
            with torch.cuda.stream(copy_s):
                q_.copy_(h_q, non_blocking=True)
                k_.copy_(h_k, non_blocking=True)
                v_.copy_(h_v, non_blocking=True)
                in_ready[si].record(copy_s)
            stream.wait_event(in_ready[si])
            if i >= 2:
                stream.wait_event(d2h_done[si])  # step i - 2's output of this set is on the host
            run_layers(si)
            layers_done[si].record(stream)
            evictions += bool(mgr.evict_after_decode())
            back_s.wait_event(layers_done[si])
            with torch.cuda.stream(back_s):
                h_out.copy_(o_, non_blocking=True)
                d2h_done[si].record(back_s)

        def steps_to_interval_start() -> int:
            return (budget + interval - int(mgr.row_seq_lens[0])) % interval

        for _ in range(steps_to_interval_start() + interval):  # warm-up: into steady state, one whole interval
            step(False)
        torch.cuda.synchronize()
        sampler = ClockSampler(0)
        sampler.start()
        n_int = max(1, round(args.steps / interval))
        steps = n_int * interval
        ev0 = evictions
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step(False)
        e1.record(stream)
        torch.cuda.synchronize()
        dev_ms = e0.elapsed_time(e1)
        assert evictions - ev0 == n_int, (evictions, ev0, n_int)

        timers = KernelTimers(mgr._get_provider(hq))
        timers.enabled = True
        for _ in range(3):  # eager, event pairs around the scored attention launch
            mgr.prepare_decode_static(seqs, ids, pos, slot_mapping, ctx, req)
            layers()
            evictions += bool(mgr.evict_after_decode())
        timers.enabled = False
        torch.cuda.synchronize()
        attn_ms = timers.summary()["sparse_decode_attn"]["avg_ms"]
        attn_len = int(mgr.row_seq_lens[0])  # the rows' length at those launches (+-1)

        for _ in range(steps_to_interval_start()):
            step(True)
        torch.cuda.synchronize()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        for _ in range(steps):
            step(True)
        stream.wait_stream(back_s)  # the last step's result is on the host
        f1.record(stream)
        torch.cuda.synchronize()
        e2e_ms = f0.elapsed_time(f1)
        clocks = sampler.stop()

    ms_per_step = dev_ms / steps
    peak, peak_src = measured_peak_gbs()
    tok_bytes = 2 * hkv * HEAD_DIM * 2  # K + V rows of one token
    # per (row, layer): KV of the tokens held, their slot ids, the fp32 score row (fill + max-update + normalise)
    attn_bytes = bsz * attn_len * (tok_bytes + 4 + 3 * 4)
    mean_len = budget + (interval + 1) / 2
    step_bytes = L * bsz * mean_len * (tok_bytes + 4 + 3 * 4)
    roof = {"bound": "hbm", "kernel": "sparse_decode_attn_kernel<score2d>", "achieved": attn_bytes / (attn_ms * 1e-3) / 1e9,
            "peak": peak, "unit": "GB/s", "peak_source": peak_src, "traffic": None, "bytes_per_launch": attn_bytes,
            "avg_launch_ms": attn_ms, "tokens_per_row": attn_len}
    roof["frac"] = roof["achieved"] / peak
    roof_step = {"achieved": step_bytes / (ms_per_step * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "bytes_per_step": step_bytes}
    roof_step["frac"] = roof_step["achieved"] / peak
    launches_per_step = 3 * L + 1  # store, scored attention, normalise+accumulate per layer; score refill
    return {
        "metric": "h2o_scored_decode_tokens_per_s", "value": bsz / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": 1,
        "steps": steps, "warmup": interval, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": w["name"], "layers": L, "q_heads": hq, "kv_heads": hkv, "batch": bsz,
                   "h2o_decode_budget": budget, "h2o_decode_eviction_interval": interval, "recent_ratio": 0.5,
                   "row_tokens": f"{budget + 1}..{width}", "evictions_timed": n_int,
                   "prompt": f"{w['ctx']}-token prompt per row admitted through the provider's prefill side: {n_chunks} chunks of "
                             f"{chunk} tokens x {L} layers (KV store, window scoring of the last {window} queries, compaction "
                             f"to the budget after every chunk), {prefill_s:.1f} s of setup, not timed",
                   "l2": "inputs larger than L2 (>= 0.55 GB of KV touched per layer, 32 distinct layers)",
                   "timed": ("cuda graph replay of the layer loop" if graph is not None else "eager")
                   + " + per-step host prepare + burst eviction every interval"},
        "e2e": {"value": bsz / (e2e_ms / steps * 1e-3), "unit": UNIT, "ms_per_step": e2e_ms / steps,
                "h2d_bytes_per_step": (h_q.numel() + h_k.numel() + h_v.numel()) * 2, "d2h_bytes_per_step": h_out.numel() * 2},
        "gpu_launches": launches_per_step * steps + 1 * n_int,
        "roofline": roof, "roofline_step": roof_step, "clocks": clocks,
    }


# --------------------------------------------------------------------------------------- CPU arm
class CpuSample:
    """Bounded sample of the same workload on the host cores with the oracle port (oracle/): `rows`
    sequences of ONE layer at the workload's context / heads / budget.  tok/s is scaled to the full
    step: value = rows / (t_sample * layers)  (every layer and row costs the same)."""

    def __init__(self, w: dict, rows: int):
        import torch
        from oracle import attention as oattn, quest as oquest
        self.torch, self.oattn, self.oquest, self.w, self.rows = torch, oattn, oquest, w, rows
        torch.set_num_threads(os.cpu_count() or 1)
        self.cores = torch.get_num_threads()
        hkv, hq, ctx = w["hkv"], w["hq"], w["ctx"]
        pages_per_row = (ctx + PAGE - 1) // PAGE
        gen = torch.Generator().manual_seed(5)
        self.k = torch.randn(rows * pages_per_row * PAGE, hkv, HEAD_DIM, generator=gen).to(torch.bfloat16)
        self.v = torch.randn(rows * pages_per_row * PAGE, hkv, HEAD_DIM, generator=gen).to(torch.bfloat16)
        self.q = torch.randn(rows, hq, HEAD_DIM, generator=gen).to(torch.bfloat16)
        self.bmax, self.bmin = oquest.build_bounds_from_cache(self.k)
        perm = torch.randperm(rows * pages_per_row, generator=gen).to(torch.int32)
        self.page_table = perm.view(rows, pages_per_row).contiguous()
        self.ctx_lens = torch.tensor([ctx - 3 - b for b in range(rows)], dtype=torch.int32)
        self.req = torch.arange(rows, dtype=torch.int32)
        self.pages_per_row = pages_per_row

    def run_once(self) -> float:
        w = self.w
        t0 = time.perf_counter()
        view = self.oquest.build_decode_view_static(
            self.q, self.bmax, self.bmin, self.page_table, None, self.req, self.ctx_lens, token_budget=w["budget"],
            num_kv_heads=w["hkv"], max_context_len=w["ctx"], max_pages_per_row=self.pages_per_row, is_long_text=True)
        self.oattn.gqa_decode(self.q, self.k, self.v, view["packed_slots"], view["req_indices"], view["lens"],
                              max_len_in_batch=view["packed_slots"].shape[1])
        return time.perf_counter() - t0

    def tokens_per_s(self, seconds: float) -> float:
        return self.rows / (seconds * self.w["layers"])

    def describe(self) -> str:
        return (f"{self.rows} sequence(s) x 1 layer of {self.w['key']} (ctx {self.w['ctx']}, {self.w['hq']}q/{self.w['hkv']}kv, "
                f"budget {self.w['budget']}): page scoring + top-k + packing + sparse attention, scaled by rows/(t*layers)")


def make_cpu_sample(w: dict, rows: int):
    """The reference itself (oracle/_ref, kind "reference") when the vendored copy is there, else the oracle port."""
    from oracle import ref_arms
    why = ref_arms.reference_available()
    if why is None:
        try:
            return ref_arms.RefCpuSample(w, rows), "reference", None
        except Exception as exc:  # e.g. a dependency of the reference package is missing on this box
            why = f"importing the vendored reference failed: {type(exc).__name__}: {exc}"
    return CpuSample(w, rows), "port", why


def cpu_baseline(w: dict, budget_s: float) -> dict:
    sample, kind, why = make_cpu_sample(w, rows=2)
    sample.run_once()  # warm-up
    times, t_start = [], time.perf_counter()
    while len(times) < 3 or (time.perf_counter() - t_start < budget_s and len(times) < 50):
        times.append(sample.run_once())
    times.sort()
    med = times[len(times) // 2]
    out = {"value": sample.tokens_per_s(med), "unit": UNIT, "cores": sample.cores, "kind": kind,
           "sample": sample.describe() + f"; median of {len(times)} runs", "sample_ms": med * 1e3}
    if why:
        out["port_because"] = why
    # BASELINE config 1 runs natively on the host (no scaling): a whole decode step of the tiny model's Quest path
    try:
        w1 = dict(WORKLOADS["cfg1"], key="cfg1")
        w1["budget"] = w1["sink"] + w1["decode_keep"] + w1["recent"]
        s1, kind1, _ = make_cpu_sample(w1, rows=1)
        s1.run_once()
        t1 = sorted(s1.run_once() for _ in range(9))[4] * w1["layers"]
        out["cfg1_native"] = {"workload": w1["name"], "value": 1.0 / t1, "unit": UNIT, "ms_per_step": t1 * 1e3,
                              "kind": kind1, "cores": s1.cores}
    except Exception as exc:
        out["cfg1_native"] = {"error": f"{type(exc).__name__}: {exc}"}
    return out


def run_reference(args) -> dict:
    """`--impl reference`: the reference's CPU torch path for this hot path on the host cores (the unmodified
    package vendored under oracle/_ref; the oracle port only if that copy is unusable), same metric/config, every
    step a bounded sample.  `ms_per_step` is the full decode step the sample's rate implies (all rows, all layers);
    the measured sample time is `sample_ms`."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return {}
    w = pick_workload(args)
    if "h2o_budget" in w:
        raise SystemExit("--impl reference times the quest workloads (cfg1/cfg2/cfg3/cfg5)")
    steps, warmup = args.steps, max(args.warmup, 1)
    rows = min(2, w["bsz"])
    sample, kind, why = make_cpu_sample(w, rows=rows)
    # keep the whole run within a few minutes
    t_probe = sample.run_once()
    budget = 150.0
    steps = max(1, min(steps, int(budget / max(t_probe, 1e-3)) - warmup))
    for _ in range(warmup - 1):
        sample.run_once()
    t0 = time.perf_counter()
    for _ in range(steps):
        sample.run_once()
    per = (time.perf_counter() - t0) / steps
    value = sample.tokens_per_s(per)
    step_ms = per * 1e3 * w["layers"] * w["bsz"] / rows
    base = {"value": value, "unit": UNIT, "cores": sample.cores, "kind": kind, "sample": sample.describe()}
    if why:
        base["port_because"] = why
    return {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warmup, "ms_per_step": step_ms, "sample_ms": per * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "config": {"workload": w["name"], "layers": w["layers"], "q_heads": w["hq"], "kv_heads": w["hkv"],
                   "batch": w["bsz"], "ctx": w["ctx"], "page": PAGE, "token_budget": w["budget"],
                   "quest_skip_layers": args.skip_layers, "parallelism": "host cores",
                   "timed": f"each step = {rows} sequence(s) x 1 layer of the workload; ms_per_step = sample_ms x layers x batch / {rows}"},
        "cpu_baseline": base,
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }


def run_incumbent(args) -> dict:
    """`--impl incumbent`: the reference's own GPU path (torch selection + Triton decode kernels, oracle/_ref,
    unmodified) on this B200 — the like-for-like number the provider replaces."""
    import torch
    from oracle import ref_arms
    why = ref_arms.reference_available()
    if why is not None:
        return {"impl": "incumbent", "unavailable": why}
    if args.gpus != 1:
        return {"impl": "incumbent", "unavailable": "single-GPU arm"}
    torch.cuda.set_device(0)
    w = pick_workload(args)
    try:
        return ref_arms.run_incumbent(w, steps=max(1, min(args.steps, 5)), warmup=args.warmup)
    except Exception as exc:
        return {"impl": "incumbent", "unavailable": f"{type(exc).__name__}: {exc}"[:400]}


def incumbent_subprocess(key: str, args) -> dict:
    cmd = [sys.executable, str(ROOT / "bench.py"), "--impl", "incumbent", "--workload", key, "--steps", "3", "--warmup", "3"]
    res = None
    try:
        res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        return json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    except Exception as exc:
        return {"impl": "incumbent", "unavailable": f"{type(exc).__name__}: {exc}",
                "stderr_tail": res.stderr[-400:] if res is not None else None}


def secondary_workload(key: str, args, extra=()) -> dict:
    """The same bench on another BASELINE config in a fresh process (its own KV pools), summarised."""
    cmd = [sys.executable, str(ROOT / "bench.py"), "--workload", key, "--no-cpu-baseline", "--steps", str(args.steps),
           "--warmup", str(args.warmup), *extra]
    res = None
    try:
        res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
        d = json.loads([ln for ln in res.stdout.splitlines() if ln.startswith("{")][-1])
    except Exception as exc:  # the primary line must not depend on the secondary one
        return {"error": f"{type(exc).__name__}: {exc}", "stderr_tail": res.stderr[-400:] if res is not None else None}
    out = {"workload": d["config"]["workload"], "metric": d["metric"], "value": d["value"], "unit": d["unit"],
           "ms_per_step": d["ms_per_step"], "steps": d["steps"], "e2e": d["e2e"]["value"],
           "roofline_kernel": d["roofline"]["kernel"], "roofline_frac": d["roofline"]["frac"],
           "roofline_launch_ms": d["roofline"]["avg_launch_ms"], "roofline_step_frac": d["roofline_step"]["frac"]}
    if "roofline_attention" in d:
        out["roofline_attention_frac"] = d["roofline_attention"]["frac"]
        out["attention_launch_ms"] = d["roofline_attention"]["avg_launch_ms"]
    if "roofline_score_select_pack" in d:  # the step runs one-launch layers: `roofline` is that kernel
        out["launches_per_sparse_layer"] = 1
        out["score_select_pack_frac"] = d["roofline_score_select_pack"]["frac"]
        out["score_select_pack_launch_ms"] = d["roofline_score_select_pack"]["avg_launch_ms"]
    elif "launches_per_sparse_layer" in d:
        out["launches_per_sparse_layer"] = d["launches_per_sparse_layer"]
    if "roofline_dense_layer" in d:
        out["quest_skip_layers"] = d["config"]["quest_skip_layers"]
        out["dense_layer_launch_ms"] = d["roofline_dense_layer"]["avg_launch_ms"]
        out["dense_layer_frac"] = d["roofline_dense_layer"]["frac"]
    return out


def main():
    args = parse_args()
    if args.impl == "reference":
        line = run_reference(args)
    elif args.impl == "incumbent":
        line = run_incumbent(args)
    else:
        h2o = args.workload == "cfg4"
        line = run_h2o(args) if h2o else run_b200(args)
        if line and args.gpus == 1 and args.workload == "auto":
            line["other_workloads"] = {k: secondary_workload(k, args) for k in ("cfg2", "cfg3", "cfg4")}
            line["other_workloads"]["cfg5_skip2"] = secondary_workload("cfg5", args, extra=["--skip-layers", "2"])
            # the reference's own GPU path on the same box, same workload (reported, never part of `value`)
            line["incumbent"] = incumbent_subprocess("cfg5", args)
            if "value" in line["incumbent"]:
                line["incumbent"]["speedup_of_value"] = line["value"] / line["incumbent"]["value"]
        if line and not args.no_cpu_baseline and args.gpus == 1 and not h2o:
            line["cpu_baseline"] = cpu_baseline(pick_workload(args), args.cpu_seconds)
    if line and not line.pop("_printed", False):
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
