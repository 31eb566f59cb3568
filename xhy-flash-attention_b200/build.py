"""In-tree nvcc build of libflashattn_b200.so for sm_100a.

``python xhy-flash-attention_b200/build.py`` (or ``__graft_entry__.build()``) compiles every ``csrc/*.cu`` with
``-gencode arch=compute_100a,code=sm_100a -lineinfo`` and links ``lib/libflashattn_b200.so``.  Objects are cached
under ``build/`` keyed by a hash of the source, the headers and the flags, so rebuilds only touch what changed.
nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import concurrent.futures as cf
import glob
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
ROOT = os.path.dirname(HERE)
OUT_DIR = os.path.join(HERE, "lib")
OBJ_DIR = os.path.join(HERE, "build")
LIB = os.path.join(OUT_DIR, "libflashattn_b200.so")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
    "--expt-relaxed-constexpr", "-Xcompiler", "-fPIC",
    "-I", os.path.join(ROOT, "include"), "-I", CSRC,
]


def _hash(paths, extra=""):
    h = hashlib.sha256(extra.encode())
    for p in sorted(paths):
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()[:16]


def _compile(src, headers, verbose):
    name = os.path.splitext(os.path.basename(src))[0]
    key = _hash([src] + headers, " ".join(FLAGS))
    obj = os.path.join(OBJ_DIR, f"{name}.{key}.o")
    if os.path.exists(obj):
        return obj, False, ""
    for old in glob.glob(os.path.join(OBJ_DIR, f"{name}.*.o")):
        os.remove(old)
    cmd = [NVCC] + FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    return obj, True, r.stderr


def build(verbose: bool = False, force: bool = False) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    os.makedirs(OBJ_DIR, exist_ok=True)
    if force:
        for old in glob.glob(os.path.join(OBJ_DIR, "*.o")):
            os.remove(old)
    srcs = sorted(glob.glob(os.path.join(CSRC, "*.cu")))
    headers = sorted(glob.glob(os.path.join(CSRC, "*.h")) + glob.glob(os.path.join(CSRC, "*.cuh")) +
                     glob.glob(os.path.join(ROOT, "include", "*.h")))
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, headers, verbose), srcs))
    objs = [r[0] for r in results]
    rebuilt = any(r[1] for r in results)
    if verbose:
        for r in results:
            if r[2]:
                print(r[2], file=sys.stderr)
    if rebuilt or not os.path.exists(LIB):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lcudart", "-Xlinker", "--no-undefined"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
