"""Builds libfsb200.so in-tree with nvcc for sm_100a (no JIT cache, no torch extension).

    python -m foundationstereo_b200.build [--force]
"""
from __future__ import annotations

import glob
import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
LIB = os.path.join(PKG, "libfsb200.so")
STAMP = os.path.join(PKG, "csrc", ".build_stamp")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-shared",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
    "-cudart", "static",
]


def _sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _fingerprint() -> str:
    h = hashlib.sha256()
    files = _sources() + sorted(glob.glob(os.path.join(CSRC, "*.cuh")))
    files.append(os.path.join(os.path.dirname(PKG), "include", "fsb200.h"))
    for f in files:
        h.update(f.encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def is_stale() -> bool:
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as fh:
        return fh.read().strip() != _fingerprint()


def _compile_one(nvcc: str, src: str, obj: str, deps_hash: str):
    """One .cu -> .o (cached by the hash of the source, the shared headers and the flags)."""
    tag = obj + ".hash"
    h = hashlib.sha256(deps_hash.encode())
    with open(src, "rb") as fh:
        h.update(fh.read())
    digest = h.hexdigest()
    if os.path.exists(obj) and os.path.exists(tag):
        with open(tag) as fh:
            if fh.read().strip() == digest:
                return 0, f"[cached] {os.path.basename(src)}\n"
    flags = [f for f in NVCC_FLAGS if f != "-shared"]
    cmd = [nvcc] + flags + ["-c", "-o", obj, src]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode == 0:
        with open(tag, "w") as fh:
            fh.write(digest)
    return res.returncode, " ".join(cmd) + "\n" + res.stdout + res.stderr


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile csrc/*.cu (one object per file, in parallel, cached under csrc/build/) and link
    libfsb200.so; returns the library path."""
    if not force and not is_stale():
        return LIB
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: libfsb200.so cannot be built")
    objdir = os.path.join(CSRC, "build")
    os.makedirs(objdir, exist_ok=True)
    if force:
        for f in glob.glob(os.path.join(objdir, "*.hash")):
            os.remove(f)
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for f in sorted(glob.glob(os.path.join(CSRC, "*.cuh"))) + [os.path.join(os.path.dirname(PKG), "include", "fsb200.h")]:
        with open(f, "rb") as fh:
            h.update(fh.read())
    deps_hash = h.hexdigest()
    from concurrent.futures import ThreadPoolExecutor
    srcs = _sources()
    objs = [os.path.join(objdir, os.path.basename(s)[:-3] + ".o") for s in srcs]
    with ThreadPoolExecutor(max_workers=min(len(srcs), os.cpu_count() or 1)) as pool:
        results = list(pool.map(lambda so: _compile_one(nvcc, so[0], so[1], deps_hash), zip(srcs, objs)))
    log = "".join(r[1] for r in results)
    failed = any(r[0] != 0 for r in results)
    if not failed:
        cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-cudart", "static", "-o", LIB] + objs
        res = subprocess.run(cmd, capture_output=True, text=True)
        log += " ".join(cmd) + "\n" + res.stdout + res.stderr
        failed = res.returncode != 0
    with open(os.path.join(CSRC, ".build_log.txt"), "w") as fh:
        fh.write(log)
    if failed:
        raise RuntimeError("nvcc failed:\n" + log[-8000:])
    if verbose:
        print(log)
    with open(STAMP, "w") as fh:
        fh.write(_fingerprint())
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
