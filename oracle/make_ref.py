"""Stage the UNMODIFIED reference package for the GPU box.  Test/measurement infrastructure, not product code.

    python oracle/make_ref.py            (build container only: needs /root/reference)

Copies the reference's Python package `autovc/` (source files only, byte for byte) from `/root/reference` into
`oracle/_ref/autovc/`.  `oracle/_ref/` is git-ignored (no reference source enters the history) but NOT gpurun-ignored, so
it travels to the GPU box with the snapshot the same way the built `.so` does.  `bench.py --impl reference`, the
`cpu_baseline` leg and the `gpu_eager_baseline` leg import the reference `AutoEncoder` from there through
`oracle/ref_loader.py` and run it through its own public API - nothing is edited, wrapped or re-implemented.
A manifest with the SHA-256 of every copied file is written next to the copy; `ref_loader.verify()` re-checks it.

`__graft_entry__.build()` calls `stage()` whenever `/root/reference` exists.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = "/root/reference"
DST = os.path.join(HERE, "_ref")
PACKAGE = "autovc"


def _sha(path: str) -> str:
    h = hashlib.sha256()
    with open(path, "rb") as f:
        h.update(f.read())
    return h.hexdigest()


def stage(src: str = SRC, dst: str = DST) -> str | None:
    """Copy `src/autovc/**/*.py` to `dst/autovc/`; returns the destination or None when the reference is absent."""
    pkg = os.path.join(src, PACKAGE)
    if not os.path.isdir(pkg):
        return None
    out = os.path.join(dst, PACKAGE)
    if os.path.isdir(out):
        shutil.rmtree(out)
    manifest = {}
    for root, _dirs, files in os.walk(pkg):
        for f in sorted(files):
            if not f.endswith(".py"):
                continue
            s = os.path.join(root, f)
            rel = os.path.relpath(s, src)
            d = os.path.join(dst, rel)
            os.makedirs(os.path.dirname(d), exist_ok=True)
            shutil.copyfile(s, d)
            manifest[rel] = _sha(d)
    with open(os.path.join(dst, "MANIFEST.json"), "w") as f:
        json.dump({"source": src, "files": manifest}, f, indent=1, sort_keys=True)
    return dst


if __name__ == "__main__":
    r = stage()
    if r is None:
        print(f"{SRC} not found: nothing staged", file=sys.stderr)
        sys.exit(1)
    print(r)
