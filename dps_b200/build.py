"""Build the native library in-tree: csrc/*.cu -> dps_b200/_native/libdps_b200.so (sm_100a only).

nvcc cross-compiles without a GPU, so this runs on the CPU build box; the resulting .so is
git-ignored but travels to the GPU box with the repo snapshot.  ``python -m dps_b200.build``.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor
from typing import List

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CSRC = os.path.join(ROOT, "csrc")
INCLUDE = os.path.join(ROOT, "include")
OUT_DIR = os.path.join(ROOT, "dps_b200", "_native")
OBJ_DIR = os.path.join(ROOT, "build", "obj")
LIB_PATH = os.path.join(OUT_DIR, "libdps_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
    "-I", INCLUDE, "-I", CSRC,
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    return "nvcc"


def sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path: str) -> str:
    h = hashlib.sha256()
    for p in [path] + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))) \
            + sorted(os.path.join(INCLUDE, f) for f in os.listdir(INCLUDE)):
        with open(p, "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def _compile_one(src: str, verbose: bool, ptxas_v: bool) -> str:
    os.makedirs(OBJ_DIR, exist_ok=True)
    base = os.path.splitext(os.path.basename(src))[0]
    obj = os.path.join(OBJ_DIR, base + ".o")
    stamp = obj + ".sha"
    dig = _digest(src)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if ptxas_v else []) + ["-c", src, "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if ptxas_v:
        log = os.path.join(OBJ_DIR, base + ".ptxas.txt")
        with open(log, "w") as f:
            f.write(r.stderr)
    with open(stamp, "w") as f:
        f.write(dig)
    return obj


def build(verbose: bool = False, force: bool = False, ptxas_v: bool = True) -> str:
    os.makedirs(OUT_DIR, exist_ok=True)
    if force and os.path.isdir(OBJ_DIR):
        for f in os.listdir(OBJ_DIR):
            os.remove(os.path.join(OBJ_DIR, f))
    srcs = sources()
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(s, verbose, ptxas_v), srcs))
    newest = max(os.path.getmtime(o) for o in objs)
    if not os.path.exists(LIB_PATH) or os.path.getmtime(LIB_PATH) < newest or force:
        cmd = [nvcc(), "-shared", "-o", LIB_PATH] + objs + ["-gencode", "arch=compute_100a,code=sm_100a",
                                                           "-Xcompiler", "-fPIC", "-lcudart_static", "-ldl", "-lrt",
                                                           "-lpthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB_PATH


if __name__ == "__main__":
    path = build(verbose="-v" in sys.argv, force="--force" in sys.argv)
    print(path)
