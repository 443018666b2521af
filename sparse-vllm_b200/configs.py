"""Quest / H2O runtime parameters with the reference's names, defaults and validation.

Follows configs/groups.py:46-66 (defaults), configs/sparse.py:33-98 (Quest budget derivation,
H2O validation incl. the ``(budget + interval) % 64 == 0`` rule) and configs/runtime_params.py:35-72,
126-198 (public names: ``sparse_method``, ``sink_keep_tokens``, ``decode_keep_tokens``,
``recent_keep_tokens``; legacy/internal keys such as ``quest_token_budget`` are rejected).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any

from .method_registry import normalize_sparse_method, require_provider_method

PAGE = 16

# configs/runtime_params.py:35-72 (the keys relevant to this path)
_LEGACY_RUNTIME_KEYS = {
    "model_cls": "sparse_method",
    "vllm_sparse_method": "sparse_method",
    "num_top_tokens": "decode_keep_tokens",
    "num_sink_tokens": "sink_keep_tokens",
    "num_recent_tokens": "recent_keep_tokens",
    "tail_token_size": "recent_keep_tokens",
    "quest_token_budget": ("removed; Sparse-vLLM QuEST derives it from sink_keep_tokens + "
                           "decode_keep_tokens + recent_keep_tokens"),
}
_PUBLIC_TO_NATIVE = {"sink_keep_tokens": "num_sink_tokens", "recent_keep_tokens": "num_recent_tokens"}


def derive_quest_token_budget(*, sink_keep_tokens: int, decode_keep_tokens: int, recent_keep_tokens: int) -> int:
    """configs/sparse.py:40-58."""
    for name, value in (("num_sink_tokens", sink_keep_tokens), ("decode_keep_tokens", decode_keep_tokens),
                        ("num_recent_tokens", recent_keep_tokens)):
        if isinstance(value, bool) or not isinstance(value, int) or value < 0:
            raise ValueError(f"QuEST {name} must be a non-negative integer, got {value!r}.")
    budget = sink_keep_tokens + decode_keep_tokens + recent_keep_tokens
    if budget <= 0:
        raise ValueError("QuEST derived token budget must be > 0: num_sink_tokens + decode_keep_tokens + "
                         f"num_recent_tokens = {budget}.")
    return budget


def quest_page_budget(token_budget: int, page_size: int = PAGE, max_pages: int | None = None) -> tuple[int, int]:
    """(page_budget_base, prev_budget) of quest.py:1846-1859: ``max(3, budget // page)`` pages of which
    the last is always the newest page."""
    base = max(3, int(token_budget) // int(page_size))
    prev = base - 1 if max_pages is None else min(base - 1, int(max_pages) - 1)
    return base, prev


@dataclass(kw_only=True)
class SparseMethodConfig:
    """The Quest / H2O fields of the reference's ``SparseMethodConfig`` (configs/groups.py:42-66)."""

    vllm_sparse_method: str = ""
    num_sink_tokens: int = 64
    num_recent_tokens: int = 512
    decode_keep_tokens: int = 4096
    quest_chunk_size: int = 16
    quest_token_budget: int = field(init=False, default=0)
    quest_skip_layers: int = 2
    h2o_decode_budget: int = 4096
    h2o_decode_eviction_interval: int = 128
    h2o_prefill_budget: int = 8192
    h2o_recent_ratio: float = 0.5
    h2o_prefill_score_window: int = 128

    def __post_init__(self):
        self.vllm_sparse_method = normalize_sparse_method(self.vllm_sparse_method)
        self._normalize_quest()
        if self.vllm_sparse_method == "h2o":
            self._normalize_h2o()

    def _normalize_quest(self):  # configs/sparse.py:33-60
        if self.quest_chunk_size <= 0:
            raise ValueError("quest_chunk_size must be > 0")
        self.quest_token_budget = 0
        if self.vllm_sparse_method == "quest":
            if self.quest_chunk_size != PAGE:
                raise ValueError(f"the B200 provider supports quest_chunk_size={PAGE} only, got {self.quest_chunk_size}.")
            self.quest_token_budget = derive_quest_token_budget(
                sink_keep_tokens=self.num_sink_tokens, decode_keep_tokens=self.decode_keep_tokens,
                recent_keep_tokens=self.num_recent_tokens)
        if self.quest_skip_layers < 0:
            raise ValueError("quest_skip_layers must be >= 0")

    def _normalize_h2o(self):  # configs/sparse.py:62-98
        for name in ("h2o_decode_budget", "h2o_decode_eviction_interval"):
            value = getattr(self, name)
            if isinstance(value, bool) or not isinstance(value, int) or value <= 0:
                raise ValueError(f"{name} must be a positive integer, got {value!r}.")
        if self.h2o_prefill_budget < self.h2o_decode_budget:
            raise ValueError("h2o_prefill_budget must be >= h2o_decode_budget, "
                             f"got prefill={self.h2o_prefill_budget} decode={self.h2o_decode_budget}.")
        peak = self.h2o_decode_budget + self.h2o_decode_eviction_interval
        if peak % 64 != 0:
            raise ValueError("h2o_decode_budget + h2o_decode_eviction_interval must be divisible by 64 for the "
                             f"scored decode kernel, got {self.h2o_decode_budget} + "
                             f"{self.h2o_decode_eviction_interval} = {peak}.")
        self.h2o_recent_ratio = float(self.h2o_recent_ratio)
        if not 0.0 < self.h2o_recent_ratio < 1.0:
            raise ValueError(f"h2o_recent_ratio must be in (0, 1), got {self.h2o_recent_ratio}.")
        if not 1 <= int(self.h2o_prefill_score_window) <= 128:  # configs/sparse.py:94-98 (probability mode)
            raise ValueError("h2o_prefill_score_window must be in [1, 128] because the prefill score kernel supports at "
                             f"most 128 query tokens, got {self.h2o_prefill_score_window}.")

    @property
    def h2o_decode_capacity(self) -> int:
        """decode_cuda_graph_context_capacity of H2O: budget + interval (h2o.py:241-254)."""
        return self.h2o_decode_budget + self.h2o_decode_eviction_interval


def normalize_runtime_params(params: dict[str, Any] | None) -> dict[str, Any]:
    """Public kwargs (``LLM(model, sparse_method=..., **kw)``) -> native field names.

    configs/runtime_params.py:126-198: legacy keys raise, ``sparse_method`` becomes
    ``vllm_sparse_method``, the public token-budget names map to the native ones."""
    out = dict(params or {})
    found = sorted(k for k in out if k in _LEGACY_RUNTIME_KEYS)
    if found:
        details = ", ".join(f"`{k}` -> `{_LEGACY_RUNTIME_KEYS[k]}`" for k in found)
        raise ValueError("Legacy runtime parameter names are no longer accepted. "
                         f"Use the new semantic names instead: {details}.")
    method = out.pop("sparse_method", None)
    if method is not None:
        out["vllm_sparse_method"] = normalize_sparse_method(str(method))
    for public, native in _PUBLIC_TO_NATIVE.items():
        if public in out:
            out[native] = out.pop(public)
    value = out.get("decode_keep_tokens")
    if isinstance(value, float) and value <= 1.0:
        raise ValueError(f"Sparse-vLLM `decode_keep_tokens` must be an explicit token count, got ratio-style value {value!r}.")
    return out


def make_sparse_config(**public_kwargs) -> SparseMethodConfig:
    """``sparse_method="quest"|"h2o"`` + public kwargs -> validated config; unknown keys are errors
    (engine/llm_engine.py:217-230)."""
    native = normalize_runtime_params(public_kwargs)
    known = set(SparseMethodConfig.__dataclass_fields__) - {"quest_token_budget"}
    unknown = sorted(set(native) - known)
    if unknown:
        raise ValueError(f"Unknown runtime parameter(s) for the Quest/H2O provider: {unknown}.")
    cfg = SparseMethodConfig(**native)
    require_provider_method(cfg.vllm_sparse_method)
    return cfg
