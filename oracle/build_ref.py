#!/usr/bin/env python
"""Vendor the UNMODIFIED reference package into oracle/_ref/ (git-ignored; it travels to the GPU box with the
repo snapshot like the built .so files).

    python oracle/build_ref.py          # needs /root/reference (build container only)

The reference is pure Python (its GPU kernels are Triton), so "building" it is copying
/root/reference/src/sparsevllm byte for byte; nothing is patched.  `bench.py --impl reference` times its CPU torch
path (QuestCacheManager._build_decode_view_static, quest.py:1833-1913) and `bench.py --impl incumbent` its
torch + Triton GPU path from this copy; the product package never imports it.  A manifest with the file count
and a digest of the copied tree is written next to it so a stale or partial copy is detected.
"""
from __future__ import annotations

import hashlib
import shutil
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
SRC = Path("/root/reference/src/sparsevllm")
DST = ROOT / "oracle" / "_ref" / "sparsevllm"


def tree_digest(root: Path) -> tuple[int, str]:
    h = hashlib.sha256()
    n = 0
    for path in sorted(root.rglob("*.py")):
        h.update(str(path.relative_to(root)).encode())
        h.update(path.read_bytes())
        n += 1
    return n, h.hexdigest()


def build(force: bool = False) -> Path | None:
    if not SRC.exists():
        return DST if DST.exists() else None
    n_src, d_src = tree_digest(SRC)
    manifest = DST.parent / "MANIFEST"
    if not force and DST.exists() and manifest.exists() and manifest.read_text().split() == [str(n_src), d_src]:
        return DST
    if DST.exists():
        shutil.rmtree(DST)
    DST.parent.mkdir(parents=True, exist_ok=True)
    shutil.copytree(SRC, DST, ignore=shutil.ignore_patterns("__pycache__", "*.pyc"))
    n_dst, d_dst = tree_digest(DST)
    assert (n_dst, d_dst) == (n_src, d_src), "copy differs from the source tree"
    manifest.write_text(f"{n_dst} {d_dst}\n")
    return DST


if __name__ == "__main__":
    out = build(force="--force" in sys.argv)
    print(out if out else "no /root/reference and no oracle/_ref copy")
