"""sparse-vllm_b200: B200 (sm_100a) provider of Sparse-vLLM's decode-time sparse-attention hot path.

Quest page selection + sparse paged GQA decode attention, and the H2O scored decode, as
hand-written CUDA behind a C ABI (``include/svb200.h`` / ``lib/libsvb200.so``), wrapped by a
host-side mirror of the reference's cache-manager hooks, attention backend and operator-provider
registry.  There is no CPU, Triton or library fallback anywhere in this package.

The directory name contains a hyphen, so the package is imported under the module name
``sparse_vllm_b200`` via ``svb200_import.py`` at the repository root.
"""
from . import _native, ops  # noqa: F401

__all__ = ["_native", "ops"]
