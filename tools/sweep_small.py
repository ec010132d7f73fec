"""Small-batch sweep (the way the reference CALLS search: nq = 1 serve, 32 eval chunk, 128 recsys): achieved HBM GB/s
of the main kernel = N*ld*sizeof(T) / kernel time (library CUDA events), per nq.

    python tools/sweep_small.py [--config c2] [--n N] [--nqs 1,2,4,8,32,128] [--k K]
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--n", type=int, default=0)
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--nqs", default="1,2,4,8,32,128")
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--store", default="", help="override the storage dtype (f32 | bf16 | f16)")
    ap.add_argument("--flags", type=int, default=0, help="NXF_SEARCH_* flags (4 = force K1, 32 = force K2T, 2 = force K2)")
    a = ap.parse_args()
    import torch
    from nexus_b200 import FlatIndex
    Q, N, D, k, store, norm, desc = bench.CONFIGS[a.config]
    N, k = a.n or N, a.k or k
    store = a.store or store
    dev = torch.device("cuda", 0)
    idx = FlatIndex(D, "ip", store, device=0)
    idx.reserve(N)
    idx.search_flags = a.flags
    bench.build_shard(torch, idx, 0, N, D, store, norm, dev)
    g = torch.Generator(device=dev)
    g.manual_seed(2002)
    q_all = torch.randn((256, D), generator=g, device=dev)
    if norm:
        q_all = torch.nn.functional.normalize(q_all, dim=-1)
    peaks = bench.load_peaks()
    esz = 4 if store == "f32" else 2
    ld = (D + 63) // 64 * 64
    for nq in [int(x) for x in a.nqs.split(",")]:
        q = q_all[:nq].contiguous()
        for _ in range(3):
            idx.search(q, k)
        torch.cuda.synchronize()
        idx.set_profiling(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            idx.search(q, k)
        e1.record()
        torch.cuda.synchronize()
        st = idx.stats()
        idx.set_profiling(False)
        kms = st["main_kernel_ms"] / max(1, st["main_kernel_count"])
        total = e0.elapsed_time(e1) / a.iters
        # bytes the kernel streams per element: the tensor paths read 16-bit rows (fp32 storage: its bf16 plane), K2 / KX the rows as stored
        esz_k = 2 if (store != "f32" or st["last_path_name"] in ("tstream", "tensor")) else 4
        gbs = N * ld * esz_k / (kms * 1e-3) / 1e9
        print(json.dumps({"config": a.config, "n": N, "dim": D, "store": store, "k": k, "nq": nq, "path": st["last_path_name"], "streamed_bytes_per_element": esz_k, "kernel_ms": round(kms, 4),
                          "call_ms": round(total, 4), "GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / peaks["hbm_gbs"], 3),
                          "qps": round(nq / (total * 1e-3), 1), "grid": st["last_grid"], "chunks": st["last_chunks"]}), flush=True)


if __name__ == "__main__":
    main()
