"""Sparse decode-attention operator family and its sm_100a provider.

Shape of the family follows the reference's ``operators/decode_attention.py`` (spec dataclass 12-27,
provider classes 30-128, ``prepare_*_op`` 148-158): the reference ships an H100-only tuned launch and a
generic Triton default; this package registers the hand-written CUDA provider, gated on compute
capability (10, 0).  There is deliberately no lower-priority generic provider: off-B200 resolution
fails loudly instead of falling back.
"""
from __future__ import annotations

from dataclasses import dataclass

import torch

from .. import ops
from .registry import DeviceCaps, OpRegistry, OpResolver, PlatformEnum, SupportResult, current_device_caps

HEAD_DIM = 128
PAGE = 16


@dataclass(frozen=True)
class SparseDecodeAttentionSpec:
    num_query_heads: int
    num_kv_heads: int
    head_dim: int
    activation_dtype: torch.dtype
    page_size: int = 1           # quest: 16; h2o (token-granular slots): 1
    sparse_method: str = ""

    def __post_init__(self) -> None:
        if self.num_query_heads <= 0 or self.num_kv_heads <= 0:
            raise ValueError("Decode attention head counts must be positive.")
        if self.num_query_heads % self.num_kv_heads:
            raise ValueError("Decode query heads must be divisible by KV heads.")
        if self.head_dim <= 0 or self.page_size <= 0:
            raise ValueError("Decode attention dimensions must be positive.")


SPARSE_DECODE_ATTENTION_REGISTRY: OpRegistry = OpRegistry("sparse decode attention")


@SPARSE_DECODE_ATTENTION_REGISTRY.register
class B200SparseDecodeProvider:
    """Quest score/select/pack + split-K paged GQA decode (+ H2O 2-D score) on sm_100a."""

    name = "b200_sm100a_sparse_decode"
    priority = 200

    @classmethod
    def supports(cls, spec: SparseDecodeAttentionSpec, caps: DeviceCaps) -> SupportResult:
        if caps.platform != PlatformEnum.CUDA or caps.compute_capability != (10, 0):
            return SupportResult.no(f"requires CUDA SM100 (B200), got {caps.platform.name} {caps.compute_capability}")
        if spec.activation_dtype != torch.bfloat16:
            return SupportResult.no(f"requires BF16 query/KV tensors, got {spec.activation_dtype}")
        if spec.head_dim != HEAD_DIM:
            return SupportResult.no(f"requires head_dim={HEAD_DIM}, got {spec.head_dim}")
        if spec.num_query_heads // spec.num_kv_heads > 16:
            return SupportResult.no(f"GQA group size {spec.num_query_heads // spec.num_kv_heads} > 16")
        if spec.sparse_method == "quest" and spec.page_size != PAGE:
            return SupportResult.no(f"quest requires page_size={PAGE}, got {spec.page_size}")
        if spec.sparse_method not in ("quest", "h2o", ""):
            return SupportResult.no(f"sparse_method={spec.sparse_method!r} is not served by this provider")
        return SupportResult.yes()

    def __init__(self, sm_count: int | None = None) -> None:
        self._sm_count = sm_count
        self._stats: dict[str, dict[str, int]] = {}

    # ---- launch policy --------------------------------------------------------------------
    @property
    def sm_count(self) -> int:
        if self._sm_count is None:
            self._sm_count = ops.sm_count()
        return self._sm_count

    def heads_per_cta(self, num_kv_heads: int) -> int:
        """KV heads staged together by one CTA: 0 = the library's choice (up to 4 that divide the local head count: a
        token's K row is then HPC*256 contiguous bytes; experiment knob 9 overrides it).  Load balance does not depend
        on it: the kernel cuts the launch's 16-token pages into one equal range per SM (stream-K), whatever the batch
        and row lengths are."""
        del num_kv_heads
        return 0

    def workspace(self, batch: int, num_q_heads: int, num_kv_heads: int, device) -> torch.Tensor:
        """Zeroed tickets + per-SM partial records (replaces the reference's mid_o / mid_o_logexpsum)."""
        return ops.attn_workspace(batch, num_q_heads, num_kv_heads, device)

    # ---- ops ------------------------------------------------------------------------------
    def _count(self, path: str, key: str = "launches") -> None:
        d = self._stats.setdefault(path, {})
        d[key] = d.get(key, 0) + 1

    def quest_select(self, q, bounds_layer, unit_table, page_table, req_rows, context_lens, *, n_cols, num_kv_heads,
                     prev_budget, scores, packed_slots, sparse_lens, tickets, selected_pages=None,
                     dense_token_table=None, token_budget=0, page_budget_base=0, out_width=0,
                     store: ops.DecodeStore | None = None) -> None:
        """``store``: the step's new K/V rows, not yet in the cache.  The long-text launch writes them itself;
        the two-launch (mixed) path stores them first with the stand-alone kernel."""
        if dense_token_table is None:
            ops.quest_score_select_pack(q, bounds_layer, unit_table, page_table, req_rows, context_lens, n_cols, num_kv_heads,
                                        prev_budget, scores, packed_slots, sparse_lens, tickets, selected_pages,
                                        store=store)
            self._count("quest_store_score_select_pack_fused" if store is not None else "quest_score_select_pack_fused")
        else:
            if store is not None:
                ops.store_kv_minmax(store.key, store.value, store.k_cache, store.v_cache, store.slot_mapping,
                                    bounds_layer if store.refresh_bounds else None, store.page_rec)
                self._count("store_kv_minmax")
            ops.quest_page_score(q, bounds_layer, unit_table, req_rows, context_lens, n_cols, num_kv_heads, scores,
                                 sched=tickets)
            ops.quest_select_pack(scores, n_cols, page_table, req_rows, context_lens, prev_budget, packed_slots,
                                  sparse_lens, selected_pages, dense_token_table, token_budget, page_budget_base,
                                  out_width)
            self._count("quest_page_score")
            self._count("quest_select_pack_mixed")

    def quest_decode_layer(self, q, bounds_layer, unit_table, page_table, req_rows, context_lens, *, n_cols,
                           num_kv_heads, prev_budget, scores, packed_slots, sparse_lens, tickets, k_cache, v_cache,
                           out, workspace, store: ops.DecodeStore | None = None, comm=None, gathered=None) -> torch.Tensor:
        """One launch for a sparse long-text decode layer: (store +) score + select + pack + sparse attention."""
        ops.quest_decode_layer(q, bounds_layer, unit_table, page_table, req_rows, context_lens, n_cols, num_kv_heads,
                               prev_budget, scores, packed_slots, sparse_lens, tickets, k_cache, v_cache, out=out,
                               workspace=workspace, store=store, comm=comm, gathered=gathered)
        self._count("quest_decode_layer_one_launch" + ("_push" if comm is not None and comm.world > 1 else ""))
        return out

    def decode(self, q, k_cache, v_cache, slot_table, req_rows, seq_lens, *, out, workspace, max_len: int,
               attn_score=None, comm=None, gathered=None) -> torch.Tensor:
        """``comm`` / ``gathered``: tensor-parallel push epilogue (ops.sparse_decode_attn)."""
        ops.sparse_decode_attn(q, k_cache, v_cache, slot_table, req_rows, seq_lens, out=out, workspace=workspace,
                               max_len=max_len, attn_score=attn_score,
                               heads_per_cta=self.heads_per_cta(int(k_cache.shape[1])), comm=comm, gathered=gathered)
        self._count(("sparse_decode_attn_score2d" if attn_score is not None else "sparse_decode_attn")
                    + ("_tp_push" if comm is not None and comm.world > 1 else ""))
        return out

    def runtime_kernel_stats(self) -> dict[str, object]:
        return {"kernel_paths": {k: dict(v) for k, v in self._stats.items()}, "fallback_reasons": {}}


class PreparedSparseDecodeAttentionOp:
    def __init__(self, spec: SparseDecodeAttentionSpec, provider: B200SparseDecodeProvider) -> None:
        self.spec = spec
        self.provider = provider

    @property
    def name(self) -> str:
        return self.provider.name


def prepare_sparse_decode_attention_op(spec: SparseDecodeAttentionSpec, *, device_index: int | None = None,
                                       caps: DeviceCaps | None = None) -> PreparedSparseDecodeAttentionOp:
    """Resolve the provider for the current device (operators/decode_attention.py:148-158)."""
    if caps is None:
        caps = current_device_caps(device_index)
    provider = OpResolver(SPARSE_DECODE_ATTENTION_REGISTRY).resolve(spec, caps).provider
    return PreparedSparseDecodeAttentionOp(spec, provider)
