"""The C-ABI library loads and exports every symbol include/svb200.h declares (no GPU compute)."""
from __future__ import annotations

import re
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent


def _header_functions() -> list[str]:
    text = (ROOT / "include" / "svb200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(svb200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_minimum_set():
    names = _header_functions()
    for required in ("svb200_quest_page_minmax", "svb200_quest_score_select_pack", "svb200_sparse_decode_attn",
                     "svb200_h2o_normalize_accumulate", "svb200_h2o_select_compact", "svb200_store_kv_minmax",
                     "svb200_workspace_bytes", "svb200_sparse_decode_attn_tp", "svb200_allreduce_sum_bf16"):
        assert required in names


def test_library_exports_every_declared_symbol(native_lib):
    for name in _header_functions():
        assert hasattr(native_lib, name), f"libsvb200.so does not export {name}"


def test_binding_table_matches_header(native_lib):
    from sparse_vllm_b200 import _native
    assert sorted(_native.EXPORTED_SYMBOLS) == _header_functions()
    assert native_lib.svb200_abi_version() == _native.ABI_VERSION


def test_binding_arity_matches_header():
    """ctypes argtypes count == parameter count of each prototype."""
    from sparse_vllm_b200 import _native
    text = (ROOT / "include" / "svb200.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    for name, (_, argtypes) in _native._SIGNATURES.items():
        m = re.search(r"\b" + name + r"\s*\((.*?)\)\s*;", text, flags=re.S)
        assert m, name
        params = m.group(1).strip()
        n = 0 if params in ("", "void") else params.count(",") + 1
        assert n == len(argtypes), f"{name}: header has {n} parameters, binding {len(argtypes)}"


def test_validation_errors_need_no_gpu(native_lib):
    """Argument validation happens before any CUDA call: bad arguments return EINVAL with a message."""
    rc = native_lib.svb200_quest_page_minmax(None, None, 8, None, None, 4, 8, 1024, None)
    assert rc == 1
    assert b"null pointer" in native_lib.svb200_last_error()
    # tickets (256-byte rounded) + two (output + lse) fp32 partial records per SM (148 without a device)
    assert native_lib.svb200_workspace_bytes(32, 28, 4) == 512 + 148 * 2 * 28 * (128 + 4) * 4
    rc = native_lib.svb200_h2o_select_compact(1, 1, 1, 1, None, None, None, 2, None, 100, 200, 10, None)
    assert rc == 1 and b"kv_len" in native_lib.svb200_last_error()


def test_product_package_has_no_oracle_or_cpu_fallback():
    """The shipped package never imports oracle/ and refuses CPU tensors."""
    import pytest
    import torch
    from sparse_vllm_b200 import ops
    pkg = ROOT / "sparse-vllm_b200"
    for path in pkg.rglob("*.py"):
        src = path.read_text()
        assert "import oracle" not in src and "from oracle" not in src, path
        assert "import triton" not in src, path
    k = torch.zeros(32, 1, 128, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="no CPU path"):
        ops.quest_page_minmax(k, torch.zeros(ops.bounds_shape(2, 1), dtype=torch.bfloat16), torch.zeros(2, dtype=torch.int32),
                              torch.zeros(1, dtype=torch.int32))
