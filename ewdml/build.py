"""Build the native library in-tree with nvcc for sm_100a.

``python -m ewdml.build`` compiles every source under ``csrc/`` to an object file (in parallel)
and links ``ewdml/_C/libewdml.so``.  The .so is git-ignored but travels to the GPU box with the
repo snapshot.  Objects are rebuilt only when a source or header is newer.
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
OUT_DIR = os.path.join(ROOT, "ewdml", "_C")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB = os.path.join(OUT_DIR, "libewdml.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "-Xcompiler", "-Wno-unused-function",
              "--expt-relaxed-constexpr", "-I", os.path.join(ROOT, "include")]


def nvcc() -> str:
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def sources():
    pats = ["**/*.cu", "**/*.cpp"]
    out = []
    for p in pats:
        out += glob.glob(os.path.join(CSRC, p), recursive=True)
    return sorted(out)


def _headers_mtime() -> float:
    hs = glob.glob(os.path.join(CSRC, "**/*.h"), recursive=True) + glob.glob(os.path.join(CSRC, "**/*.cuh"), recursive=True) \
        + glob.glob(os.path.join(ROOT, "include", "*.h"))
    return max([os.path.getmtime(h) for h in hs] + [0.0])


def _compile(src: str, verbose: bool, ptxas_v: bool) -> str:
    rel = os.path.relpath(src, CSRC).replace(os.sep, "_")
    obj = os.path.join(OBJ_DIR, rel + ".o")
    if os.path.exists(obj) and os.path.getmtime(obj) > max(os.path.getmtime(src), _headers_mtime()):
        return obj
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_v else []) + ["-x", "cu", "-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if ptxas_v or verbose:
        sys.stderr.write(r.stderr)
    return obj


def build(verbose: bool = False, force: bool = False, ptxas_v: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for f in glob.glob(os.path.join(OBJ_DIR, "*.o")):
            os.remove(f)
    srcs = sources()
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile(s, verbose, ptxas_v), srcs))
    if (not os.path.exists(LIB)) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        cmd = [nvcc(), "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                         "-cudart", "static", "-lpthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    path = build(verbose="-v" in sys.argv, force="-f" in sys.argv, ptxas_v="--ptxas" in sys.argv)
    print(path)
