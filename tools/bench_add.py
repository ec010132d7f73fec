"""Index build (`index()` = astype(float32) + faiss add, utils.py:165-189) timed on its own: host rows -> HBM shard.

    python tools/bench_add.py [--n 4000000] [--d 768] [--src f32] [--store bf16]
Reports rows/s and host GB/s for (a) a pageable numpy array (what the reference hands over), (b) a pinned torch tensor,
(c) a CUDA tensor (the device-resident encode -> index hand-off, N1), with NXF_ADD_THREADS = 1 and 8 for the host cases.
Host path = threads copy into double-buffered pinned staging, async H2D, convert / re-pitch kernel; bound by PCIe.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(a, threads):
    os.environ["NXF_ADD_THREADS"] = str(threads)
    import numpy as np
    import torch
    from nexus_b200 import FlatIndex
    npdt = {"f32": np.float32, "f16": np.float16}[a.src]
    rng = np.random.default_rng(0)
    x = np.empty((a.n, a.d), dtype=npdt)
    for r0 in range(0, a.n, 1 << 18):
        x[r0:r0 + (1 << 18)] = rng.standard_normal((min(1 << 18, a.n - r0), a.d), dtype=np.float32).astype(npdt)
    gb = x.nbytes / 1e9
    out = {"threads": threads, "n": a.n, "d": a.d, "src": a.src, "store": a.store, "host_gb": round(gb, 3)}

    def timed(src, label):
        best = None
        for _ in range(a.reps):
            idx = FlatIndex(a.d, "ip", a.store, device=0)
            idx.reserve(a.n)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            idx.add(src)
            torch.cuda.synchronize()
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
            assert idx.ntotal == a.n
            del idx
        out[label] = {"s": round(best, 4), "rows_per_s": round(a.n / best, 1), "src_gb_per_s": round(gb / best, 2)}

    timed(x, "numpy_pageable")
    xt = torch.from_numpy(x).pin_memory()
    timed(xt, "torch_pinned")
    if threads == a.threads[-1]:
        timed(xt.cuda(), "cuda_tensor")
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=4_000_000)
    ap.add_argument("--d", type=int, default=768)
    ap.add_argument("--src", default="f32")
    ap.add_argument("--store", default="bf16")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--threads", type=int, nargs="+", default=[1, 8])
    ap.add_argument("--one", type=int, default=None, help=argparse.SUPPRESS)
    a = ap.parse_args()
    if a.one is not None:
        run(a, a.one)
        return
    for t in a.threads:   # one process per setting: the library reads NXF_ADD_THREADS per call, but a fresh pinned pool keeps runs independent
        subprocess.run([sys.executable, __file__, "--n", str(a.n), "--d", str(a.d), "--src", a.src, "--store", a.store, "--reps", str(a.reps),
                        "--threads", *[str(x) for x in a.threads], "--one", str(t)], check=True)


if __name__ == "__main__":
    main()
