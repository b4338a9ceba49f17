"""Build libsplatloc_b200.so in-tree with nvcc for sm_100a (no torch headers).

    python -m splatloc_b200.build [--force] [--verbose]

Each .cu under csrc/ is compiled to an object in csrc/_build/ (in parallel, only
when out of date) and linked into splatloc_b200/libsplatloc_b200.so.  nvcc
cross-compiles without a GPU, so this runs in the CPU-only container; the built
.so travels to the GPU box with the repo snapshot.
"""
from __future__ import annotations

import argparse
import concurrent.futures as cf
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
BUILD = os.path.join(CSRC, "_build")
LIB = os.path.join(HERE, "libsplatloc_b200.so")

NVCC_FLAGS = [
    "-O3", "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    nvcc = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found; cannot build libsplatloc_b200.so")
    return nvcc


def _sources() -> list[str]:
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _headers_mtime() -> float:
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "splatloc_b200.h"))
    return max(os.path.getmtime(h) for h in hs)


def _compile(src: str, verbose: bool, extra: list[str]) -> str:
    obj = os.path.join(BUILD, src[:-3] + ".o")
    # SL_NVCC_EXTRA: extra flags for tuning experiments, e.g. "-DSL_BLEND_CTA_WARPS=8"
    extra = [*extra, *os.environ.get("SL_NVCC_EXTRA", "").split()]
    cmd = [_nvcc(), *NVCC_FLAGS, *extra, "-c", os.path.join(CSRC, src), "-o", obj]
    if verbose:
        print(" ".join(cmd), flush=True)
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError(f"nvcc failed for {src}:\n{r.stdout}\n{r.stderr}")
    if verbose and (r.stdout or r.stderr):
        print(r.stdout, r.stderr, flush=True)
    return obj


def build(force: bool = False, verbose: bool = False, ptxas_info: bool = False) -> str:
    os.makedirs(BUILD, exist_ok=True)
    srcs = _sources()
    hdr_t = _headers_mtime()
    extra = ["-Xptxas", "-v"] if ptxas_info else []
    todo = []
    for s in srcs:
        obj = os.path.join(BUILD, s[:-3] + ".o")
        if (force or ptxas_info or not os.path.exists(obj)
                or os.path.getmtime(obj) < max(os.path.getmtime(os.path.join(CSRC, s)), hdr_t)):
            todo.append(s)
    if todo:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(todo))) as ex:
            list(ex.map(lambda s: _compile(s, verbose or ptxas_info, extra), todo))
    objs = [os.path.join(BUILD, s[:-3] + ".o") for s in srcs]
    if todo or not os.path.exists(LIB) or os.path.getmtime(LIB) < max(os.path.getmtime(o) for o in objs):
        cmd = [_nvcc(), "-shared", "-gencode", "arch=compute_100a,code=sm_100a", *objs, "-o", LIB]
        if verbose:
            print(" ".join(cmd), flush=True)
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    ap.add_argument("--ptxas-info", action="store_true", help="recompile with -Xptxas -v (registers, spills, smem)")
    a = ap.parse_args()
    print(build(a.force, a.verbose, a.ptxas_info))
