"""Import helper: registers the directory ``sparse-vllm_b200/`` as the package ``sparse_vllm_b200``.

``import svb200_import`` (from the repo root or with the repo root on sys.path) makes
``import sparse_vllm_b200`` work; the hyphenated directory name is mandated by the repo layout.
"""
from __future__ import annotations

import importlib.util
import sys
from pathlib import Path

PKG_NAME = "sparse_vllm_b200"
PKG_DIR = Path(__file__).resolve().parent / "sparse-vllm_b200"


def load():
    if PKG_NAME in sys.modules:
        return sys.modules[PKG_NAME]
    spec = importlib.util.spec_from_file_location(
        PKG_NAME, PKG_DIR / "__init__.py", submodule_search_locations=[str(PKG_DIR)])
    module = importlib.util.module_from_spec(spec)
    sys.modules[PKG_NAME] = module
    try:
        spec.loader.exec_module(module)
    except BaseException:
        sys.modules.pop(PKG_NAME, None)
        raise
    return module


load()
