"""Two ranks on two GPUs (skipped on a single-GPU box; run under `gpurun --gpus 2`): every rank's Quest selection
and attention on ITS head shard against the oracle evaluated on that shard (selection is TP-local in the
reference, CONTEXT.md:27-29), the attention epilogue's NVLink push against the concatenated shards in
`[B, Hq, D]` order, and the one-shot NVLink all-reduce against NCCL and an fp32 sum."""
from __future__ import annotations

import os
import socket
import sys
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
pytestmark = pytest.mark.gpu
PAGE = 16


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank: int, world: int, port: int) -> None:
    sys.path.insert(0, str(ROOT))
    import svb200_import  # noqa: F401
    import torch.distributed as dist
    from oracle import attention as oattn
    from oracle import cases, quest as oquest
    from sparse_vllm_b200 import configs, ops
    from sparse_vllm_b200.operators.all_reduce import (AllReduceOpSpec, B200NvlinkOneShotAllReduceProvider,
                                                       TorchDistributedAllReduceProvider, prepare_all_reduce_op)
    from sparse_vllm_b200.parallel import PeerComm, shard_heads

    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    try:
        # ---------------------------------------------------------------- Quest decode layer, KV heads sharded
        cfg = dict(seed=61, bsz=3, hq=16, hkv=4, ctx_lens=[1024, 1003, 770], max_context_len=1024, token_budget=256)
        inp = cases.make_quest_inputs(**cfg)
        sh = shard_heads(cfg["hq"], cfg["hkv"], world, rank)
        hq, hkv, bsz = sh.q_heads, sh.kv_heads, cfg["bsz"]
        k_rows = [k[:, sh.kv_slice()].contiguous() for k in inp["k_rows"]]
        v_rows = [v[:, sh.kv_slice()].contiguous() for v in inp["v_rows"]]
        q = inp["q"][:, sh.q_slice()].contiguous()
        num_slots = inp["num_pages"] * PAGE
        kc = torch.zeros(num_slots, hkv, 128, dtype=torch.bfloat16, device=dev)
        vc = torch.zeros_like(kc)
        ut, rec, num_chunks = cases.chunk_maps(inp["page_table"], inp["num_pages"], seed=5)
        ut, rec = ut.to(dev), rec.to(dev)
        bounds = ops.alloc_bounds(num_chunks, hkv, dev)
        for b in range(bsz):
            slots = inp["slot_rows"][b].to(dev)
            ops.store_kv_minmax(k_rows[b].to(dev), v_rows[b].to(dev), kc, vc, slots)
            ops.quest_page_minmax(kc, bounds, rec, slots)
        max_pages = inp["max_pages_per_row"]
        n_cols = max_pages - 1
        _, prev = configs.quest_page_budget(cfg["token_budget"], PAGE, max_pages)
        req, ctx = inp["req_indices"].to(dev), inp["context_lens"].to(dev)
        scores = torch.empty(bsz, ops.score_row_stride(n_cols), dtype=torch.bfloat16, device=dev)
        packed = torch.zeros(bsz, (prev + 1) * PAGE, dtype=torch.int32, device=dev)
        lens = torch.zeros(bsz, dtype=torch.int32, device=dev)
        sel = torch.zeros(bsz, prev + 1, dtype=torch.int32, device=dev)
        ops.quest_score_select_pack(q.to(dev), bounds, ut, inp["page_table"].to(dev), req, ctx, n_cols, hkv, prev, scores,
                                    packed, lens, ops.sched_scratch(bsz, dev), sel)
        # oracle on this rank's shard
        k_cache = torch.zeros(num_slots, hkv, 128, dtype=torch.bfloat16)
        v_cache = torch.zeros_like(k_cache)
        for k, v, slots in zip(k_rows, v_rows, inp["slot_rows"]):
            k_cache.index_copy_(0, slots.long(), k)
            v_cache.index_copy_(0, slots.long(), v)
        mx, mn = oquest.build_bounds_from_cache(k_cache)
        view = oquest.build_decode_view_static(q, mx, mn, inp["page_table"], inp["token_table"], inp["req_indices"],
                                               inp["context_lens"], token_budget=cfg["token_budget"], num_kv_heads=hkv,
                                               max_context_len=cfg["max_context_len"], max_pages_per_row=max_pages)
        torch.cuda.synchronize()
        assert torch.equal(lens.cpu(), view["lens"])
        n_valid = (inp["context_lens"] + PAGE - 1) // PAGE - 1
        for b in range(bsz):
            c = oquest.classify_selection_ulp(view["scores"][b], view["selected_pages"][b, :prev].tolist(),
                                              sel[b, :prev].cpu().tolist(), int(n_valid[b]), ulp_slack=1)
            assert c["bad"] == 0, f"rank {rank} row {b}: {c['detail']}"
        # attention with the push epilogue: every rank ends up with [B, Hq_total, D]
        # twice: the flag-in-data protocol (staging behind the sync block), then fences + flags (no staging)
        comm_flags = PeerComm(2 * bsz * cfg["hq"] * 128 * 2, ll_bytes=0)
        comm_ll = PeerComm(2 * bsz * cfg["hq"] * 128 * 2)
        assert comm_ll.ll_bytes > 0 and comm_flags.ll_bytes == 0
        out = torch.empty(bsz, hq, 128, dtype=torch.bfloat16, device=dev)
        arange = torch.arange(bsz, dtype=torch.int32, device=dev)
        ws = ops.attn_workspace(bsz, hq, hkv, dev)
        for it in range(10):  # alternating buffers, epochs / flags only grow
            comm = comm_flags if it >= 6 else comm_ll
            gathered = [comm.view(i * bsz * cfg["hq"] * 256, (bsz, cfg["hq"], 128), torch.bfloat16) for i in range(2)]
            gathered[it & 1].fill_(float("nan"))
            dist.barrier()
            torch.cuda.synchronize()
            ops.sparse_decode_attn(q.to(dev), kc, vc, packed, arange, lens, out=out, workspace=ws, comm=comm,
                                   gathered=gathered[it & 1])
            torch.cuda.synchronize()  # the launch completes only after every rank's rows arrived
            ref = oattn.gqa_decode(q, k_cache, v_cache, packed.cpu(), torch.arange(bsz, dtype=torch.int32), lens.cpu(),
                                   max_len_in_batch=packed.shape[1])
            torch.testing.assert_close(out.cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
            shards = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(shards, out)
            want = torch.cat(shards, dim=1)  # heads of rank 0, then rank 1: the [B, Hq, D] o_proj reads
            assert torch.equal(gathered[it & 1], want), f"rank {rank} iteration {it}: pushed rows differ from the shards"
        # the one-launch layer (score + select + pack + attention) with the same epilogue
        gathered = [comm_ll.view(i * bsz * cfg["hq"] * 256, (bsz, cfg["hq"], 128), torch.bfloat16) for i in range(2)]
        packed2, lens2 = torch.zeros_like(packed), torch.zeros_like(lens)
        for it in range(2):
            gathered[it & 1].fill_(float("nan"))
            dist.barrier()
            torch.cuda.synchronize()
            ops.quest_decode_layer(q.to(dev), bounds, ut, inp["page_table"].to(dev), req, ctx, n_cols, hkv, prev, scores,
                                   packed2, lens2, ops.sched_scratch(bsz, dev), kc, vc, out=out, workspace=ws,
                                   comm=comm_ll, gathered=gathered[it & 1])
            torch.cuda.synchronize()
            assert torch.equal(packed2, packed) and torch.equal(lens2, lens)
            shards = [torch.empty_like(out) for _ in range(world)]
            dist.all_gather(shards, out)
            assert torch.equal(gathered[it & 1], torch.cat(shards, dim=1)), f"rank {rank}: one-launch layer, push {it}"
            torch.testing.assert_close(out.cpu().float(), ref.float(), atol=2e-2, rtol=1e-2)
        # ---------------------------------------------------------------- all-reduce provider family
        rows, hidden = 16, 5120
        spec = AllReduceOpSpec(world_size=world, ranks=tuple(range(world)), max_rows=rows, hidden_size=hidden,
                               dtype=torch.bfloat16, cuda_graph=True, backend="nccl")
        op = prepare_all_reduce_op(spec, group=None, rank=rank)
        assert isinstance(op.provider, B200NvlinkOneShotAllReduceProvider), op.name
        nccl = prepare_all_reduce_op(spec, group=dist.group.WORLD, rank=rank, provider=TorchDistributedAllReduceProvider())
        g = torch.Generator(device=dev).manual_seed(100 + rank)
        for it in range(5):
            x = torch.randn(rows, hidden, device=dev, generator=g).to(torch.bfloat16)
            if it % 2:  # write straight into the staging buffer (no copy inside run)
                stage = op.provider.input_buffer(rows, hidden)
                stage.copy_(x)
                got = op.run(stage)
            else:
                got = op.run(x.clone())
            parts = [torch.empty_like(x) for _ in range(world)]
            dist.all_gather(parts, x)
            want = torch.stack(parts).float().sum(dim=0).to(torch.bfloat16)  # fp32 sum in rank order, one rounding
            assert torch.equal(got, want), f"rank {rank} all-reduce {it}"
            ref2 = nccl.run(x.clone())
            torch.testing.assert_close(got.float(), ref2.float(), atol=2e-2, rtol=2e-2)
        op.close()
        comm_ll.close()
        comm_flags.close()
    finally:
        dist.destroy_process_group()


def test_tp2_quest_layer_push_and_allreduce():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(2, _free_port()), nprocs=2, join=True)
