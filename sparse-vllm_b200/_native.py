"""ctypes binding of libsvb200.so (include/svb200.h) — the only door to the CUDA kernels.

There is deliberately NO fallback: if the shared library is missing or a call returns a
non-zero code, a RuntimeError is raised (the reference is fail-fast too,
AGENTS.md:51-57 / kernels/triton/gqa_flash_decoding_stage1.py:347-352).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_float, c_int32, c_int64, c_void_p
from pathlib import Path

import torch

_LIB_PATH = Path(__file__).resolve().parent / "lib" / "libsvb200.so"
ABI_VERSION = 6

MAX_PEERS = 8


class Comm(ctypes.Structure):
    """svb200_comm_t: peer-mapped data buffers and sync blocks of every rank."""
    _fields_ = [("world", c_int32), ("rank", c_int32), ("peer_buf", c_void_p * MAX_PEERS),
                ("peer_sync", c_void_p * MAX_PEERS), ("peer_ll", c_void_p * MAX_PEERS), ("ll_bytes", c_int64)]


# name -> (restype, argtypes); mirrors include/svb200.h one to one.
_P = c_void_p
_ATTN_ARGS = [_P, c_int64, _P, _P, c_int64, _P, c_int64, c_int32, _P, _P, _P, c_int64, _P, c_int64,
              c_int32, c_int32, c_int32, c_int32, _P, c_int64, c_float]
_SIGNATURES = {
    "svb200_abi_version": (c_int32, []),
    "svb200_last_error": (c_char_p, []),
    "svb200_sm_count": (c_int32, []),
    "svb200_set_max_ctas": (c_int32, [c_int32]),
    "svb200_set_tuning": (c_int32, [c_int32, c_int32]),
    "svb200_sched_scratch_ints": (c_int64, [c_int32]),
    "svb200_debug_times": (None, [_P]),
    "svb200_workspace_bytes": (c_int64, [c_int32, c_int32, c_int32]),
    "svb200_quest_page_minmax": (c_int32, [_P, _P, c_int64, _P, _P, c_int32, c_int32, c_int64, _P]),
    "svb200_quest_decode_page_complete": (
        c_int32, [_P, _P, c_int64, _P, _P, c_int64, _P, _P, c_int32, c_int32, c_int32, c_int64, c_int64, _P]),
    "svb200_quest_page_score": (
        c_int32, [_P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, c_int32, c_int32, c_int32,
                  _P, _P]),
    "svb200_quest_select_pack": (
        c_int32, [_P, c_int64, c_int32, _P, c_int64, _P, _P, c_int32, c_int32, _P, c_int64, _P, _P, _P,
                  c_int64, c_int32, c_int32, c_int32, _P]),
    "svb200_quest_score_select_pack": (
        c_int32, [_P, c_int64, _P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, c_int32,
                  c_int32, c_int32, c_int32, _P, c_int64, _P, _P, _P, _P]),
    "svb200_quest_score_select_pack_static": (  # the one above without `tickets`: the scratch lives in the library
        c_int32, [_P, c_int64, _P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, c_int32,
                  c_int32, c_int32, c_int32, _P, c_int64, _P, _P, _P]),
    "svb200_quest_store_score_select_pack": (
        c_int32, [_P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, _P,
                  _P, c_int64, _P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, c_int32,
                  c_int32, c_int32, c_int32, _P, c_int64, _P, _P, _P, _P]),
    "svb200_quest_decode_layer": (
        c_int32, [_P, c_int64, _P, c_int64, _P, c_int32, _P, _P, _P, c_int64,
                  _P, c_int64, _P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, c_int64, c_int32, c_int32,
                  c_int32, c_int32, c_int32, _P, c_int64, _P, _P, _P,
                  _P, c_int64, _P, c_int64, c_float, POINTER(Comm), c_int64, c_int64, _P]),
    "svb200_sparse_decode_attn": (c_int32, [*_ATTN_ARGS, _P]),
    "svb200_sparse_decode_attn_tp": (c_int32, [*_ATTN_ARGS, POINTER(Comm), c_int64, c_int64, _P]),
    "svb200_allreduce_sum_bf16": (c_int32, [POINTER(Comm), c_int64, c_int64, _P, _P]),
    "svb200_symm_alloc": (c_int32, [c_int64, POINTER(c_void_p), _P]),
    "svb200_symm_free": (c_int32, [_P]),
    "svb200_ipc_open": (c_int32, [_P, POINTER(c_void_p)]),
    "svb200_ipc_close": (c_int32, [_P]),
    "svb200_h2o_normalize_accumulate": (
        c_int32, [_P, c_int64, c_int32, c_int32, c_float, _P, c_int64, _P, _P, c_float, _P]),
    "svb200_h2o_prepare_decode": (
        c_int32, [_P, _P, c_int64, _P, c_int64, c_int64, _P, c_int64, _P, _P, _P, _P, _P, c_int32, c_int32, _P]),
    "svb200_prefill_attn": (
        c_int32, [_P, c_int64, _P, _P, c_int64, _P, c_int64, _P, _P, _P, _P, _P, c_int64, c_int32, c_int32, c_int32,
                  c_int32, c_float, _P]),
    "svb200_h2o_prefill_score": (
        c_int32, [_P, c_int64, _P, c_int64, _P, c_int64, _P, _P, _P, _P, _P, _P, _P, c_int64, _P, c_int64, _P,
                  c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_float, _P]),
    "svb200_h2o_select_compact": (
        c_int32, [_P, _P, _P, _P, _P, _P, _P, c_int32, _P, c_int32, c_int32, c_int32, _P]),
    "svb200_store_kv_minmax": (
        c_int32, [_P, c_int64, _P, c_int64, _P, _P, _P, c_int32, c_int32, c_int64, _P, _P, c_int64, _P]),
}
EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def library_path() -> Path:
    return _LIB_PATH


def load() -> ctypes.CDLL:
    """Load libsvb200.so once.  Raises if it is absent or has the wrong ABI."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise RuntimeError(
            f"{_LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  The B200 provider has no CPU or Triton fallback."
        )
    lib = ctypes.CDLL(str(_LIB_PATH), mode=os.RTLD_LOCAL if hasattr(os, "RTLD_LOCAL") else 0)
    for name, (restype, argtypes) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError here = header and library out of sync
        fn.restype = restype
        fn.argtypes = argtypes
    got = lib.svb200_abi_version()
    if got != ABI_VERSION:
        raise RuntimeError(f"libsvb200.so ABI version {got} != expected {ABI_VERSION}; rebuild.")
    _lib = lib
    return lib


def is_loaded() -> bool:
    return _lib is not None


def last_error() -> str:
    return (load().svb200_last_error() or b"").decode("utf-8", "replace")


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {last_error()}")


def ptr(t: torch.Tensor | None) -> int | None:
    """Device pointer of a tensor (None -> NULL)."""
    if t is None:
        return None
    return t.data_ptr()


def current_stream_ptr(device: torch.device | None = None) -> int:
    """Stream of ``device`` the launch goes to.  The library launches on the thread's CURRENT device, so a
    tensor living on another device is refused instead of launching on the wrong GPU (fail fast)."""
    if device is not None and device.index is not None and device.index != torch.cuda.current_device():
        raise RuntimeError(f"tensors live on {device} but the current CUDA device is {torch.cuda.current_device()}: "
                           "call torch.cuda.set_device(...) (one process per GPU) before using the provider.")
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(t: torch.Tensor, name: str, dtype: torch.dtype | None = None) -> None:
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (the B200 provider has no CPU path), got {t.device}.")
    if dtype is not None and t.dtype != dtype:
        raise TypeError(f"{name} must be {dtype}, got {t.dtype}.")


def sm_count() -> int:
    return int(load().svb200_sm_count())
