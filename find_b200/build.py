"""find_b200/build.py — in-tree nvcc build of libfindb200.so for sm_100a.

The library is plain CUDA behind a C ABI (include/find_b200.h): no torch headers, so it compiles in seconds
and cross-compiles on a CPU-only box.  The .so is written next to the sources (git-ignored, travels to the
GPU box with the tree).
"""
from __future__ import annotations

import hashlib
import os
import shutil
import subprocess
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(_HERE, "libfindb200.so")
_STAMP = os.path.join(_HERE, ".libfindb200.stamp")

SOURCES = ["capi.cu", "chamfer.cu", "field.cu", "field_chain.cu", "field_chain2.cu", "field_dw.cu", "step.cu", "mesh.cu", "optim.cu", "allreduce.cu", "raster.cu", "umma_selftest.cu", "diag.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "--shared", "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
    "-ccbin", "/usr/bin/g++",
]


def _nvcc() -> str:
    for c in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found: the CUDA extension cannot be built")


def _sources():
    return [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]


def _digest() -> str:
    h = hashlib.sha256()
    root = os.path.dirname(_HERE)
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [os.path.join(root, "include", "find_b200.h")]
    for f in files:
        with open(f, "rb") as fh:
            h.update(f.encode())
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _file_digest(src: str) -> str:
    """Digest of one translation unit: its source, every header under csrc/ and include/, and the flags."""
    h = hashlib.sha256()
    root = os.path.dirname(_HERE)
    deps = [src] + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    deps.append(os.path.join(root, "include", "find_b200.h"))
    for f in deps:
        with open(f, "rb") as fh:
            h.update(f.encode())
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ into libfindb200.so (no-op when sources are unchanged).  Translation units are
    compiled in parallel into _obj/*.o (git-ignored) and only re-compiled when they or a header changed."""
    from concurrent.futures import ThreadPoolExecutor
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(_STAMP):
        with open(_STAMP) as fh:
            if fh.read().strip() == dig:
                return LIB
    objdir = os.path.join(_HERE, "_obj")
    os.makedirs(objdir, exist_ok=True)
    env = dict(os.environ)
    env.pop("CC", None)
    env.pop("CXX", None)
    cflags = [f for f in NVCC_FLAGS if f != "--shared"]

    def compile_one(src: str):
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        stamp = obj + ".stamp"
        d = _file_digest(src)
        if not force and os.path.exists(obj) and os.path.exists(stamp):
            with open(stamp) as fh:
                if fh.read().strip() == d:
                    return obj, 0, ""
        cmd = [_nvcc()] + cflags + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
        if res.returncode == 0:
            with open(stamp, "w") as fh:
                fh.write(d)
        return obj, res.returncode, res.stdout

    with ThreadPoolExecutor(max_workers=min(8, os.cpu_count() or 1)) as ex:
        results = list(ex.map(compile_one, _sources()))
    log = "".join(r[2] for r in results)
    if verbose or any(r[1] for r in results):
        print(log, file=sys.stderr)
    if any(r[1] for r in results):
        raise RuntimeError("nvcc failed:\n" + log[-6000:])
    cmd = [_nvcc(), "--shared", "-Xcompiler", "-fPIC", "-ccbin", "/usr/bin/g++", "-gencode", "arch=compute_100a,code=sm_100a",
           "-o", LIB] + [r[0] for r in results]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
    if res.returncode:
        raise RuntimeError("nvcc link failed:\n" + res.stdout[-4000:])
    with open(_STAMP, "w") as fh:
        fh.write(dig)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
