"""Cache managers of the B200 provider (mirror of the reference's ``engine/cache_manager`` package)."""
from .base import (AttentionViewMeta, CacheManager, DecodeComputeView, ExplicitKVPayload, LayerBatchStates,  # noqa: F401
                   SparseSelection)
from .quest import QuestCacheManager  # noqa: F401
