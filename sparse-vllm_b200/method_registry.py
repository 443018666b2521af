"""Sparse-method names routed to the B200 provider.

Mirrors the Quest / H2O entries of the reference's ``method_registry.py`` (19-48 aliases and
canonical set, 199-209 TP decode-graph allow-list, 211-237 prefill policy) — the other methods of
the reference are not served by this provider and are rejected, not silently remapped.
"""
from __future__ import annotations

PREFILL_POLICY_ALL_CHUNKED = "all_chunked"

# method_registry.py:19-35 (only the spellings that resolve to a method this provider serves)
METHOD_ALIASES = {None: "", "": "", "vanilla": ""}

# subset of CANONICAL_SPARSE_METHODS (method_registry.py:37-48) implemented here
PROVIDER_SPARSE_METHODS = frozenset({"quest", "h2o"})
CANONICAL_SPARSE_METHODS = frozenset({"", "streamingllm", "snapkv", "h2o", "pyramidkv", "omnikv", "quest",
                                      "rkv", "skipkv", "deltakv"})

DECODE_CUDA_GRAPH_SUPPORTED_METHODS = frozenset(PROVIDER_SPARSE_METHODS)      # method_registry.py:197
TP_DECODE_CUDA_GRAPH_SUPPORTED_METHODS = frozenset(PROVIDER_SPARSE_METHODS)   # method_registry.py:199-209
PREFILL_POLICY_BY_METHOD = {"quest": PREFILL_POLICY_ALL_CHUNKED, "h2o": PREFILL_POLICY_ALL_CHUNKED}


def normalize_sparse_method(method: str | None) -> str:
    """Alias -> canonical name (method_registry.py:240-246); unknown names raise like configs/sparse.py:25-30."""
    if method is None:
        return ""
    name = str(method)
    name = METHOD_ALIASES.get(name, name)
    if name not in CANONICAL_SPARSE_METHODS:
        supported = ", ".join(repr(m) for m in sorted(CANONICAL_SPARSE_METHODS) if m)
        raise ValueError(f"Unsupported vllm_sparse_method={name!r}. Supported methods: '', {supported}.")
    return name


def provider_serves(method: str | None) -> bool:
    """True when the B200 provider implements the decode path of ``method``."""
    return normalize_sparse_method(method) in PROVIDER_SPARSE_METHODS


def require_provider_method(method: str | None) -> str:
    name = normalize_sparse_method(method)
    if name not in PROVIDER_SPARSE_METHODS:
        raise ValueError(
            f"sparse_method={name!r} is a valid Sparse-vLLM method but is not served by the B200 provider "
            f"(serves: {sorted(PROVIDER_SPARSE_METHODS)}); there is no fallback path.")
    return name
