"""Per-forward context: the fields of the reference's ``utils/context.py:1-46`` the decode path reads."""
from __future__ import annotations

from dataclasses import dataclass

import torch


@dataclass
class Context:
    is_prefill: bool = False
    is_long_text: bool = True
    decode_mid_o: torch.Tensor | None = None
    decode_mid_o_logexpsum: torch.Tensor | None = None
    decode_workspace: torch.Tensor | None = None
    decode_workspace_key: tuple | None = None


_CONTEXT = Context()


def get_context() -> Context:
    return _CONTEXT


def set_context(is_prefill: bool, *, is_long_text: bool = True) -> None:
    _CONTEXT.is_prefill = bool(is_prefill)
    _CONTEXT.is_long_text = bool(is_long_text)


def reset_context() -> None:
    global _CONTEXT
    _CONTEXT = Context()
