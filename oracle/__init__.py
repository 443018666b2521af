"""CPU oracle for the Quest / H2O decode hot path — TEST INFRASTRUCTURE ONLY.

A torch-CPU / numpy restatement of the reference algorithm (CURRENTF/Sparse-vLLM @ 18d3104;
citations are relative to its ``src/sparsevllm/``).  Only ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs
may import this package, and only as the checker / the timed CPU baseline — never from the
product package ``sparse-vllm_b200/`` (which has no CPU path at all).

Parity status: PINNED.  ``oracle/gen_golden.py`` runs the UNMODIFIED reference (imported from
``/root/reference/src`` with ``SPARSEVLLM_PLATFORM=cpu``; Triton kernels under
``TRITON_INTERPRET=1``) on seeded inputs and stores its outputs under ``tests/golden/``;
``tests/test_oracle_golden.py`` checks every function here against those files and against the
reference's own known-answer tests (tests/test_h2o_cache_manager.py:290-338,1511-1542;
tests/test_decode_cuda_graph_tp_config.py:75-127).  The reference has NO numeric test for
Quest page bounds / scores / top-k, so for those the golden files generated
from the reference code itself are the pin.
"""
from . import attention, h2o, quest  # noqa: F401
