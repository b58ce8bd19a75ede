"""In-tree build of the native library (``libb200ddp.so``) for sm_100a.

``nvcc`` cross-compiles without a GPU, so this runs on the CPU dev box; the resulting ``.so`` is
git-ignored but travels to the GPU box with the repo snapshot.  One translation unit per ``csrc/*.cu``
compiled in parallel, then one link.  Rebuilds only what changed (mtime of the source and of every
header).
"""
from __future__ import annotations

import concurrent.futures
import glob
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REPO = os.path.dirname(PKG_DIR)
CSRC = os.path.join(REPO, "csrc")
INCLUDE = os.path.join(REPO, "include")
BUILD = os.path.join(REPO, "build")
LIB_DIR = os.path.join(PKG_DIR, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "libb200ddp.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=default",
    "--expt-relaxed-constexpr",
    "-I", INCLUDE,
]


def _nvcc() -> str:
    path = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(path):
        raise RuntimeError("nvcc not found; cannot build libb200ddp.so")
    return path


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _headers_mtime() -> float:
    hs = glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.h")) + \
        glob.glob(os.path.join(INCLUDE, "*.h"))
    return max([os.path.getmtime(h) for h in hs] + [0.0])


def _compile_one(src: str, verbose: bool, ptxas_v: bool) -> str:
    obj = os.path.join(BUILD, os.path.basename(src)[:-3] + ".o")
    newest = max(os.path.getmtime(src), _headers_mtime())
    if os.path.exists(obj) and os.path.getmtime(obj) >= newest:
        return obj
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_v else []) + ["-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, res.stdout, res.stderr))
    if ptxas_v or verbose:
        sys.stderr.write(res.stderr)
    return obj


def build(force: bool = False, verbose: bool = False, ptxas_v: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(LIB_DIR, exist_ok=True)
    if force:
        for o in glob.glob(os.path.join(BUILD, "*.o")):
            os.remove(o)
    srcs = sources()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs) or 1)) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose, ptxas_v), srcs))
    if (not os.path.exists(LIB_PATH)) or any(os.path.getmtime(o) > os.path.getmtime(LIB_PATH) for o in objs):
        cmd = [_nvcc(), "-shared", "-o", LIB_PATH] + objs + [
            "-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-cudart", "static", "-ldl", "-lpthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (res.stdout, res.stderr))
    return LIB_PATH


if __name__ == "__main__":
    p = build(force="--force" in sys.argv, verbose="-v" in sys.argv, ptxas_v="--ptxas-v" in sys.argv)
    print(p)
