"""Host-side TP logic on CPU, world_size 2 over gloo: KV-head sharding + output gather reproduce the
reference semantics evaluated with N head shards (selection is TP-local, CONTEXT.md:27-29)."""
from __future__ import annotations

import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import attention as oattn
from oracle import cases, quest as oquest


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _shard_run(inp, cfg, k_cache, v_cache, shard):
    """Reference decode of one head shard: TP-local Quest selection, then sparse attention."""
    q = inp["q"][:, shard.q_slice()].contiguous()
    kc, vc = k_cache[:, shard.kv_slice()].contiguous(), v_cache[:, shard.kv_slice()].contiguous()
    mx, mn = oquest.build_bounds_from_cache(kc)
    view = oquest.build_decode_view_static(
        q, mx, mn, inp["page_table"], inp["token_table"], inp["req_indices"], inp["context_lens"],
        token_budget=cfg["token_budget"], num_kv_heads=shard.kv_heads, max_context_len=cfg["max_context_len"],
        max_pages_per_row=inp["max_pages_per_row"], is_long_text=True)
    bsz = q.shape[0]
    out = oattn.gqa_decode(q, kc, vc, view["packed_slots"], torch.arange(bsz, dtype=torch.int32), view["lens"],
                           max_len_in_batch=view["packed_slots"].shape[1])
    return out, view["selected_pages"]


def _worker(rank, world, port, ret):
    import svb200_import  # noqa: F401
    from sparse_vllm_b200 import parallel
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        cfg = cases.QUEST_CASES["quest_small_long"]
        inp = cases.make_quest_inputs(**cfg)
        k_cache, v_cache = cases.quest_fill_cache(inp, cfg["hkv"])
        shard = parallel.shard_heads(cfg["hq"], cfg["hkv"], world, rank)
        out, _ = _shard_run(inp, cfg, k_cache, v_cache, shard)
        gathered = parallel.gather_attention_output(out)
        full = parallel.heads_first(gathered)
        # every rank holds the same gathered tensor; rank 0 checks it against both shards computed locally
        if rank == 0:
            parts = []
            for r in range(world):
                o, _ = _shard_run(inp, cfg, k_cache, v_cache, parallel.shard_heads(cfg["hq"], cfg["hkv"], world, r))
                parts.append(o)
            expect = torch.cat(parts, dim=1)
            ret["equal"] = bool(torch.equal(full, expect))
            ret["shape"] = tuple(full.shape)
        chk = full.float().sum().reshape(1).clone()
        both = [torch.zeros(1), torch.zeros(1)]
        dist.all_gather(both, chk)
        ret[f"same_{rank}"] = bool(both[0] == both[1])
    finally:
        dist.destroy_process_group()


def test_tp2_gather_matches_sharded_reference():
    world, port = 2, _free_port()
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        ret = dict(ret)
    cfg = cases.QUEST_CASES["quest_small_long"]
    assert ret["equal"] and ret["shape"] == (cfg["bsz"], cfg["hq"], 128)
    assert ret["same_0"] and ret["same_1"]


def test_shard_heads_rules():
    import svb200_import  # noqa: F401
    from sparse_vllm_b200 import parallel
    s = parallel.shard_heads(64, 8, 4, 3)  # cfg5 at TP4: 2 KV heads / 16 q heads per rank
    assert (s.q_heads, s.kv_heads, s.q_begin, s.kv_begin) == (16, 2, 48, 6)
    assert [parallel.shard_heads(64, 8, tp, 0).kv_heads for tp in (1, 2, 4, 8)] == [8, 4, 2, 1]
    with pytest.raises(ValueError, match="cannot be sharded"):
        parallel.shard_heads(28, 4, 8, 0)  # TP > Hkv would need KV replication
    with pytest.raises(ValueError):
        parallel.shard_heads(30, 4, 2, 0)


def test_selection_is_tp_local():
    """Shards select on their own heads: the page sets of two shards may differ from each other and from
    the unsharded selection (the reference never exchanges indices)."""
    import svb200_import  # noqa: F401
    from sparse_vllm_b200 import parallel
    cfg = cases.QUEST_CASES["quest_small_long"]
    inp = cases.make_quest_inputs(**cfg)
    k_cache, v_cache = cases.quest_fill_cache(inp, cfg["hkv"])
    sel = [_shard_run(inp, cfg, k_cache, v_cache, parallel.shard_heads(cfg["hq"], cfg["hkv"], 2, r))[1] for r in (0, 1)]
    assert sel[0].shape == sel[1].shape
    assert torch.equal(sel[0][:, -1], sel[1][:, -1])  # the last page is always appended
