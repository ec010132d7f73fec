"""Where the end-to-end (host buffers) search spends its time beyond the device-resident one: H2D of the queries, the search,
D2H of the results, each timed alone with a host clock around a synchronised call.    python tools/e2e_breakdown.py [--config c2]"""
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
    ap.add_argument("--config", default="c2")
    ap.add_argument("--iters", type=int, default=8)
    a = ap.parse_args()
    import torch
    from nexus_b200 import FlatIndex
    Q, N, D, k, store, norm, desc = bench.CONFIGS[a.config]
    dev = torch.device("cuda", 0)
    idx = FlatIndex(D, "ip", store, device=0)
    idx.reserve(N)
    bench.build_shard(torch, idx, 0, N, D, store, norm, dev)
    g = torch.Generator(device=dev); g.manual_seed(2002)
    q = torch.nn.functional.normalize(torch.randn((Q, D), generator=g, device=dev), dim=-1)
    qh = q.cpu().pin_memory()
    od = torch.empty((Q, k), dtype=torch.float32).pin_memory()
    oi = torch.empty((Q, k), dtype=torch.int64).pin_memory()

    def t(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(a.iters):
            fn()
            torch.cuda.synchronize()
        return (time.perf_counter() - t0) / a.iters * 1e3

    res = {}
    res["search_device_sync_each_ms"] = t(lambda: idx.search(q, k))
    res["search_host_buffers_ms"] = t(lambda: idx.search(qh, k, out=(od, oi)))
    qd = torch.empty_like(q)
    res["h2d_queries_ms"] = t(lambda: qd.copy_(qh, non_blocking=True))
    Dd, Id = idx.search(q, k)
    res["d2h_results_ms"] = t(lambda: (od.copy_(Dd, non_blocking=True), oi.copy_(Id, non_blocking=True)))
    # back-to-back device searches (what bench.py's `value` times): launches of step i+1 overlap step i
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(a.iters):
        idx.search(q, k)
    e1.record()
    torch.cuda.synchronize()
    res["search_device_back_to_back_ms"] = e0.elapsed_time(e1) / a.iters
    print(json.dumps({k_: round(v, 3) for k_, v in res.items()}))


if __name__ == "__main__":
    main()
