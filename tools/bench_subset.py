"""KS (per-query candidate lists, csrc/search_subset.cu) against the HBM roofline: achieved GB/s = algorithmic bytes / CUDA-event
time, with the reference's form (per query: stack the candidates, matmul, full torch.sort -- eval_embedding.py:384-404) timed
beside it on the same GPU and the numpy oracle on the host, both on a bounded sample of the queries.

    python tools/bench_subset.py [--n 2000000] [--d 768] [--nq 4096] [--len 1000] [--k 10] [--store bf16]
Algorithmic bytes (DESIGN.md KS): total*ld*s (every listed row once) + total*8 (row ids) + nq*d*4 + nq*k*12.
The gathered rows (total*ld*s) are several times the 126 MB L2 and the lists are random, so no flush between iterations.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--n", type=int, default=2_000_000)
    ap.add_argument("--d", type=int, default=768)
    ap.add_argument("--nq", type=int, default=4096)
    ap.add_argument("--len", type=int, default=1000)
    ap.add_argument("--k", type=int, default=10)
    ap.add_argument("--store", default="bf16")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--ref-queries", type=int, default=128)
    a = ap.parse_args()
    import numpy as np
    import torch
    import oracle
    from nexus_b200 import FlatIndex
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(11)
    esz = 4 if a.store == "f32" else 2
    ld = (a.d + 63) // 64 * 64
    c = torch.randn((a.n, a.d), generator=g, device=dev, dtype=torch.float32)
    c = torch.nn.functional.normalize(c, dim=-1)
    x = torch.nn.functional.normalize(torch.randn((a.nq, a.d), generator=g, device=dev), dim=-1)
    idx = FlatIndex(a.d, "ip", a.store, device=0)
    idx.add(c)
    # ragged lists: lengths uniform in [len/2, 3*len/2], random rows
    lens = torch.randint(max(1, a.len // 2), a.len * 3 // 2 + 1, (a.nq,), generator=g, device=dev)
    offsets = torch.zeros(a.nq + 1, dtype=torch.int64, device=dev)
    offsets[1:] = torch.cumsum(lens, 0)
    total = int(offsets[-1].item())
    rows = torch.randint(0, a.n, (total,), generator=g, device=dev, dtype=torch.int64)
    alg = total * ld * esz + total * 8 + a.nq * a.d * 4 + a.nq * a.k * 12
    for _ in range(3):
        D, P = idx.search_subsets(x, offsets, rows, a.k)
    idx.set_profiling(True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        D, P = idx.search_subsets(x, offsets, rows, a.k)
    e1.record()
    torch.cuda.synchronize()
    ms_call = e0.elapsed_time(e1) / a.iters
    st = idx.stats()
    ms_kernel = st["main_kernel_ms"] / max(1, st["main_kernel_count"])
    # the reference's form on the same GPU, candidates already resident (it uploads them per query: not charged here)
    nref = min(a.ref_queries, a.nq)
    cs = c if a.store == "f32" else (c.to(torch.bfloat16 if a.store == "bf16" else torch.float16).float())
    off_h = offsets.cpu().numpy()
    rows_h = rows.cpu().numpy()

    def torch_ref():
        out = []
        for i in range(nref):
            cand = cs[rows[off_h[i]:off_h[i + 1]]]
            sim = torch.matmul(x[i].unsqueeze(0), cand.T).squeeze(0)
            _, ranked = torch.sort(sim, descending=True)
            out.append(ranked[:a.k].cpu())
        return out
    torch_ref()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    ref_ranked = torch_ref()
    torch.cuda.synchronize()
    ms_torch = (time.perf_counter() - t0) * 1e3 * a.nq / nref
    # parity on the sample (positions may differ only inside ties)
    Ph, Dh = P.cpu().numpy(), D.cpu().numpy()
    agree = 0
    for i in range(nref):
        m = min(a.k, len(ref_ranked[i]))
        agree += int(np.array_equal(ref_ranked[i].numpy()[:m], Ph[i, :m]))
    # numpy oracle on the host (bounded sample)
    cs_h = cs[rows[: off_h[min(nref, 32)]]].cpu().numpy()
    xh = x[:32].cpu().numpy()
    t0 = time.perf_counter()
    o = 0
    for i in range(min(nref, 32)):
        ln = off_h[i + 1] - off_h[i]
        sc = oracle.mmeb_similarity(xh[i:i + 1], cs_h[o:o + ln])[0]
        np.argsort(-sc, kind="stable")
        o += ln
    ms_numpy = (time.perf_counter() - t0) * 1e3 * a.nq / min(nref, 32)
    peaks = bench.load_peaks()
    hbm = peaks["hbm_gbs"]
    gbps = alg / (ms_kernel * 1e-3) / 1e9
    print(json.dumps({
        "kernel": "subset_search_kernel", "n": a.n, "d": a.d, "nq": a.nq, "mean_list": a.len, "total_listed_rows": total, "k": a.k,
        "store": a.store, "algorithmic_bytes": alg, "ms_kernel": round(ms_kernel, 4), "ms_call": round(ms_call, 4),
        "achieved_gbps": round(gbps, 1), "hbm_peak_gbps": hbm, "frac_of_hbm_peak": round(gbps / hbm, 4) if hbm else None,
        "lists_per_s": round(a.nq / (ms_call * 1e-3), 1),
        "torch_loop_ms_scaled": round(ms_torch, 1), "numpy_oracle_ms_scaled": round(ms_numpy, 1),
        "sample_rankings_identical_to_torch": f"{agree}/{nref}",
    }))


if __name__ == "__main__":
    main()
