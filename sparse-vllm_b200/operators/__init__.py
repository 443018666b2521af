"""Operator-provider registry of the B200 path (mirror of the reference's ``operators/`` API)."""
from .registry import (DeviceCaps, OpRegistry, OpResolver, PlatformEnum, ResolvedProvider, SupportResult,  # noqa: F401
                       current_device_caps, operator_runtime_stats, record_operator_binding)
from .decode_attention import (B200SparseDecodeProvider, SPARSE_DECODE_ATTENTION_REGISTRY,  # noqa: F401
                               SparseDecodeAttentionSpec, prepare_sparse_decode_attention_op)
