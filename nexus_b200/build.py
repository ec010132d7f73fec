"""Build libnexus_flat.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m nexus_b200.build [--force] [--verbose]

The shared library is the product's only compute path; there is no JIT cache and no CPU fallback.
"""
from __future__ import annotations

import argparse
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")
LIB = os.path.join(HERE, "libnexus_flat.so")
OBJ_DIR = os.path.join(HERE, "_build")

SOURCES = ["api.cu", "convert.cu", "merge.cu", "pool.cu", "search_exact.cu", "search_stream.cu", "search_stream_tc.cu", "search_subset.cu", "search_tc.cu"]
HEADERS = ["common.cuh", "select.cuh", "sink.cuh", "params.cuh", "block_sort.cuh"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O2",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def _newest_input() -> float:
    paths = [os.path.join(CSRC, f) for f in SOURCES + HEADERS] + [os.path.join(INCLUDE, "nexus_flat.h"), __file__]
    return max(os.path.getmtime(p) for p in paths)


def needs_build() -> bool:
    return not os.path.exists(LIB) or os.path.getmtime(LIB) < _newest_input()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    nvcc = _nvcc()
    os.makedirs(OBJ_DIR, exist_ok=True)
    hdr_mtime = max(os.path.getmtime(os.path.join(CSRC, h)) for h in HEADERS)
    hdr_mtime = max(hdr_mtime, os.path.getmtime(os.path.join(INCLUDE, "nexus_flat.h")), os.path.getmtime(__file__))

    def compile_one(src: str):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ_DIR, src.replace(".cu", ".o"))
        if not force and os.path.exists(o) and os.path.getmtime(o) >= max(os.path.getmtime(s), hdr_mtime):
            return src, 0, ""
        cmd = [nvcc, *NVCC_FLAGS, *os.environ.get("NXF_EXTRA_NVCC_FLAGS", "").split(), "-I", INCLUDE, "-c", s, "-o", o]
        r = subprocess.run(cmd, capture_output=True, text=True)
        return src, r.returncode, r.stdout + r.stderr

    with ThreadPoolExecutor(max_workers=min(8, len(SOURCES))) as ex:
        results = list(ex.map(compile_one, SOURCES))
    log = []
    failed = False
    for src, rc, out in results:
        if out:
            log.append(f"==== {src} ====\n{out}")
        failed |= rc != 0
    with open(os.path.join(OBJ_DIR, "build.log"), "w") as f:
        f.write("\n".join(log))
    if failed:
        sys.stderr.write("\n".join(log) + "\n")
        raise RuntimeError("nvcc failed; see " + os.path.join(OBJ_DIR, "build.log"))
    if verbose:
        print("\n".join(log))
    objs = [os.path.join(OBJ_DIR, s.replace(".cu", ".o")) for s in SOURCES]
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB, *objs, "-cudart", "static"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("link failed")
    return LIB


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--force", action="store_true")
    ap.add_argument("--verbose", action="store_true")
    a = ap.parse_args()
    print(build(force=a.force, verbose=a.verbose))
