#!/usr/bin/env python
"""Carry the (pure-Python) reference to the GPU box: oracle/_ref/.

    python tools/make_ref.py            # in the build container, where /root/reference exists

The reference has no native sources to compile; what `oracle/_ref` holds for this repo is
the UNMODIFIED Python of the reference's model (`core/`, `Utils.py`, `depth_anything/`,
the vendored `dinov2/` package), copied file by file from where it lies under
/root/reference.  `oracle/_ref/` is git-ignored (nothing of the reference enters the history)
but not gpurun-ignored, so it travels to the GPU box exactly like the built libfsb200.so.
There it is TEST INFRASTRUCTURE: `oracle/refmodel.py` imports it as the end-to-end oracle
(stock `FoundationStereo.forward` vs the same object after `fsb.install(model)`) and
`bench.py`'s CPU legs time its hot-path functions (`cpu_baseline.kind = "reference"`).
Nothing under foundationstereo_b200/ reads it.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.environ.get("FS_REFERENCE", "/root/reference")
DST = os.path.join(ROOT, "oracle", "_ref")

# (path relative to the reference root, recurse?)  -- Python sources only
WANT = [("core", True), ("Utils.py", False), ("depth_anything", True), ("dinov2/hubconf.py", False),
        ("dinov2/dinov2", True)]


def _py_files(path: str):
    if os.path.isfile(path):
        yield path
        return
    for d, _, files in os.walk(path):
        if "__pycache__" in d:
            continue
        for f in sorted(files):
            if f.endswith(".py"):
                yield os.path.join(d, f)


def make_ref(src: str = SRC, dst: str = DST, verbose: bool = True) -> str | None:
    """Returns the destination, or None when the reference checkout is not present (GPU box)."""
    if not os.path.isdir(os.path.join(src, "core")):
        return None
    if os.path.isdir(dst):
        shutil.rmtree(dst)
    manifest = {}
    for rel, _ in WANT:
        for f in _py_files(os.path.join(src, rel)):
            r = os.path.relpath(f, src)
            out = os.path.join(dst, r)
            os.makedirs(os.path.dirname(out), exist_ok=True)
            shutil.copyfile(f, out)
            with open(f, "rb") as fh:
                manifest[r] = hashlib.sha256(fh.read()).hexdigest()
    with open(os.path.join(dst, "MANIFEST.json"), "w") as fh:
        json.dump({"source": src, "files": manifest}, fh, indent=1, sort_keys=True)
    if verbose:
        print(f"oracle/_ref: {len(manifest)} unmodified reference files copied from {src}")
    return dst


if __name__ == "__main__":
    sys.exit(0 if make_ref() else 1)
