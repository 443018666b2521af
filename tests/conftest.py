"""Shared fixtures.  `-m "not gpu"` = oracle vs golden vectors, host logic, C-ABI symbol checks
(no GPU compute); `-m gpu` = parity of the CUDA path (through the C ABI) against the oracle."""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

import svb200_import  # noqa: E402,F401  (registers sparse_vllm_b200)

GOLDEN = ROOT / "tests" / "golden"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name: str):
        with np.load(GOLDEN / f"{name}.npz") as z:
            return {k: z[k] for k in z.files}
    return load


@pytest.fixture(scope="session")
def native_lib():
    """libsvb200.so, built on demand in the CPU container (nvcc cross-compiles)."""
    from sparse_vllm_b200 import _native, build
    if not _native.library_path().exists():
        build.build_native()
    return _native.load()
