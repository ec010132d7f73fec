#!/usr/bin/env python
"""bench.py -- queries/sec of exact top-k flat search (BASELINE.json's metric) on N B200s of one node.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--config c2|c1|c5|...]
    torchrun --nnodes=1 --nproc-per-node N ... bench.py --gpus N --steps K --warmup W

A "step" is one pass of the hot path over one batch of synthetic input: all Q queries of the workload
searched against the whole (row-sharded) corpus, top-k returned.  Default workload = BASELINE config 2
(MS MARCO-passage shape: 6 980 queries x 8.8M x 768-d bf16 corpus, top-100), which fits one GPU.

  value   queries/s, whole job, inputs resident in HBM when the timed region starts (device in -> device out)
  e2e     the same metric through the host-buffer C-ABI call (pinned host queries -> host results; H2D and D2H
          inside the timed region)
  roofline   dominant kernel (K1 tcgen05 GEMM + fused top-k): algorithmic FLOPs 2*Q*(N/G)*D per launch over its
          CUDA-event duration (events recorded by the library on the launching stream, nxf_set_profiling)
  cpu_baseline   the oracle's blocked numpy restatement of the reference search on a bounded sample (rank 0, N=1)

--impl reference times the reference's own CPU implementation of the path (the oracle port of
NumpyIPIndex/search: faiss is not installable here) on the host cores, bounded sample per step.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (Q, N, D, k, store dtype, normalised, description)
    "c1": (1000, 100_000, 384, 10, "f32", True, "C1: 1k q x 100k x 384 fp32 (BGE-small shape), top-10"),
    "c2": (6980, 8_800_000, 768, 100, "bf16", True, "C2: MS MARCO-passage shape, 6980 q x 8.8M x 768 bf16 corpus, top-100"),
    "c2f": (6980, 8_800_000, 768, 100, "f32", True, "C2 shape with fp32 storage (the reference's own dtype): 6980 q x 8.8M x 768 fp32, top-100"),
    "c3": (100_000, 21_000_000, 1024, 100, "bf16", True, "C3: BGE-large shape, 100k q x 21M x 1024 bf16, top-100"),
    "c4": (50_000, 5_000_000, 1536, 10, "bf16", True, "C4: MMEB-style, 50k q x 5M x 1536, top-10"),
    "c5": (65_536, 10_000_000, 64, 500, "bf16", False, "C5: two-tower recsys, 65536 q x 10M x 64, top-500"),
}
GEN_CHUNK = 1 << 18  # rows per generator call; the seed depends on the global chunk id, so the corpus is the
                     # same whatever the number of ranks


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--config", default="c2", choices=sorted(CONFIGS))
    ap.add_argument("--nq", type=int, default=0, help="override the number of queries")
    ap.add_argument("--n", type=int, default=0, help="override the number of corpus rows")
    ap.add_argument("--k", type=int, default=0)
    ap.add_argument("--dim", type=int, default=0, help="override the embedding dimension (kernel experiments; the workload name says so)")
    ap.add_argument("--store", default="", help="override the storage dtype (kernel experiments)")
    ap.add_argument("--flags", type=int, default=0, help="nxf_search flags (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-rows", type=int, default=1_000_000)
    ap.add_argument("--cpu-sample-queries", type=int, default=512)
    return ap.parse_args()


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            d = json.load(open(p))
            return {"hbm_gbs": float(d["hbm_gbs"]), "bf16_tflops": float(d["bf16_tflops"]),
                    "bf16_tflops_sustained": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])), "source": "measured"}
        except Exception:
            pass
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device_index: int):
        self.dev = device_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.dev)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._pump, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, pw, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": max(mx), "power_w_max": max(pw), "samples": len(sm),
                "reasons": sorted(reasons)}


def probe_faiss_cpu(sample_rows, sample_q, d, k, n_full, normalised):
    """if a REAL faiss is importable on this box, time IndexFlatIP exactly as the reference calls it (32-row
    query chunks, utils.py:221-224) on the same bounded sample; otherwise say so.  Never our own shim."""
    try:
        import faiss
        if getattr(faiss, "__nexus_b200__", False) or not hasattr(faiss, "IndexFlatIP"):
            raise ImportError("only the nexus_b200 shim is registered as faiss")
    except Exception as e:  # noqa: BLE001
        return {"available": False, "note": f"Faiss CPU: unavailable on this box ({type(e).__name__})"}
    import numpy as np
    import oracle
    c = oracle.round_to_bf16(oracle.make_unit_gaussian(sample_rows, d, seed=1002, normalize=normalised))
    q = oracle.make_unit_gaussian(sample_q, d, seed=2002, normalize=normalised)
    index = faiss.index_factory(d, "Flat", faiss.METRIC_INNER_PRODUCT)
    index.add(c)
    t0 = time.perf_counter()
    for i in range(0, sample_q, 32):
        index.search(np.ascontiguousarray(q[i:i + 32], dtype=np.float32), k)
    t = time.perf_counter() - t0
    return {"available": True, "value": sample_q / (t * (n_full / float(sample_rows))), "unit": "queries/s", "seconds": t,
            "call": "IndexFlatIP.search, 32-row chunks (utils.py:221-224)"}


def cpu_reference_search(sample_rows, sample_q, d, k, n_full, normalised, steps=1, warmup=0):
    """Times the oracle port of the reference CPU search (NumpyIPIndex math, blocked) on a bounded sample and
    scales linearly in N to the full corpus: value = Q_s / (t * N / N_s).  Returns (qps, seconds per step, cores)."""
    import numpy as np
    import oracle
    cores = os.cpu_count() or 1
    c = oracle.make_unit_gaussian(sample_rows, d, seed=1002, normalize=normalised)
    c = oracle.round_to_bf16(c)
    q = oracle.make_unit_gaussian(sample_q, d, seed=2002, normalize=normalised)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.blocked_ip_search(c, q, k, block_rows=65536)
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    qps = sample_q / (t * (n_full / float(sample_rows)))
    return qps, t, cores


def run_reference(a):
    """The reference arm: the reference's CPU implementation of the path (oracle port), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    Q, N, D, k, store, norm, desc = CONFIGS[a.config]
    if a.dim or a.store:
        D, store = a.dim or D, a.store or store
        desc += f" [override: dim {D}, store {store}]"
    Q, N, k = a.nq or Q, a.n or N, a.k or k
    rows = min(a.cpu_sample_rows, N)
    sq = min(a.cpu_sample_queries, Q)
    qps, t, cores = cpu_reference_search(rows, sq, D, k, N, norm, steps=a.steps, warmup=min(a.warmup, 1))
    sample = f"{sq} queries x {rows} rows x {D} (of {Q} x {N}); queries/s scaled linearly in N"
    line = {
        "impl": "reference", "metric": "queries/sec exact top-k flat search", "value": qps, "unit": "queries/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": t * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": desc, "nq": Q, "n": N, "dim": D, "k": k,
                   "note": "faiss is not installable offline; this is the oracle port of the reference's own NumpyIPIndex/search "
                           "(Nexus/evaluation/text_retrieval/utils.py:16-31,194-230), blocked so that it fits in RAM"},
        "cpu_baseline": {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port", "sample": sample,
                         "faiss_cpu": probe_faiss_cpu(rows, sq, D, k, N, norm)},
        "e2e": {"value": qps, "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


def build_shard(torch, idx, lo, hi, D, store, norm, device):
    """Synthetic corpus rows [lo, hi): N(0,1), optionally L2-normalised, rounded ONCE to the storage type
    (the stored values are the ground-truth corpus).  Generated on the device, never on the host."""
    dt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[store]
    c0, c1 = lo // GEN_CHUNK, (hi - 1) // GEN_CHUNK if hi > lo else -1
    for cid in range(c0, c1 + 1):
        g = torch.Generator(device=device)
        g.manual_seed(1002 * 1_000_003 + cid)
        x = torch.randn((GEN_CHUNK, D), generator=g, device=device, dtype=torch.float32)
        if norm:
            x = torch.nn.functional.normalize(x, dim=-1)
        r0 = cid * GEN_CHUNK
        a_, b_ = max(lo, r0) - r0, min(hi, r0 + GEN_CHUNK) - r0
        idx.add(x[a_:b_].to(dt))
        del x


def run_ours(a):
    import numpy as np
    import torch
    import torch.distributed as dist
    from nexus_b200 import FlatIndex, merge_topk, shard_bounds
    from nexus_b200 import _lib as L
    from nexus_b200.sharded import PARTITION_MIN_BYTES, all_gather_lists, exchange_partitioned

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if a.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun the way the driver does
        port = 29500 + (os.getpid() % 2000)
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={a.gpus}", "--master-addr", "127.0.0.1",
               "--master-port", str(port), os.path.abspath(__file__)] + sys.argv[1:]
        return subprocess.call(cmd)
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py: no CUDA device; the engine has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)

    Q, N, D, k, store, norm, desc = CONFIGS[a.config]
    if a.dim or a.store:
        D, store = a.dim or D, a.store or store
        desc += f" [override: dim {D}, store {store}]"
    Q, N, k = a.nq or Q, a.n or N, a.k or k
    lo, hi = shard_bounds(N, world)[rank]
    idx = FlatIndex(D, "ip", store, device=local_rank)
    idx.reserve(hi - lo)
    idx.set_id_offset(lo)
    t_build = time.perf_counter()
    build_shard(torch, idx, lo, hi, D, store, norm, device)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build

    g = torch.Generator(device=device)
    g.manual_seed(2002)
    q = torch.randn((Q, D), generator=g, device=device, dtype=torch.float32)
    if norm:
        q = torch.nn.functional.normalize(q, dim=-1)
    q_host = q.cpu().pin_memory()
    out_d = torch.empty((Q, k), dtype=torch.float32).pin_memory()
    out_i = torch.empty((Q, k), dtype=torch.int64).pin_memory()
    flags = a.flags

    # the one exchange step: all-gather of the per-shard [Q,k] lists + merge kernel; for large Q*k (C5: 393 MB per rank) the
    # query-partitioned form (all-to-all by query block, merge, all-gather of the merged blocks: 2/world of the bytes)
    partitioned = world > 2 and Q * k * 12 * world >= PARTITION_MIN_BYTES
    if os.environ.get("NXF_BENCH_EXCHANGE") in ("gather", "partition"):        # A/B of the two exchanges
        partitioned = world > 1 and os.environ["NXF_BENCH_EXCHANGE"] == "partition"

    def exchange(D_, I_):
        if partitioned:
            return exchange_partitioned(D_, I_, world, k, "ip", None, merge_topk)
        Ds, Is = all_gather_lists(D_, I_, world)
        return merge_topk(Ds, Is, k)

    def step_dev():
        D_, I_ = idx.search(q, k, flags=flags)
        if world > 1:
            D_, I_ = exchange(D_, I_)
        return D_, I_

    def step_e2e():
        if world == 1:
            return idx.search(q_host, k, flags=flags, out=(out_d, out_i))     # host buffers through the C ABI
        qd = q_host.to(device, non_blocking=True)
        D_, I_ = idx.search(qd, k, flags=flags)
        D_, I_ = exchange(D_, I_)
        if rank == 0:
            out_d.copy_(D_, non_blocking=True)
            out_i.copy_(I_, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return out_d, out_i

    def barrier():
        if world > 1:
            dist.barrier()

    def timed(fn, steps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        torch.cuda.synchronize()
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    # ---- warm-up (>= 3) + a validity check at full size against the exact fp32 kernel ----
    W = max(a.warmup, 3)
    for _ in range(W):
        Dd, Id = step_dev()
    torch.cuda.synchronize()
    check = {}
    Dl, Il = idx.search(q[:64], k, flags=flags)
    De, Ie = idx.search(q[:64], k, flags=L.SEARCH_FORCE_EXACT)
    inter = sum(len(set(Il[r].tolist()) & set(Ie[r].tolist())) for r in range(Il.shape[0]))
    check["recall_vs_exact_fp32_kernel_64q"] = inter / float(Il.numel())
    check["max_abs_score_diff_top1"] = float((Dl[:, 0] - De[:, 0]).abs().max().item())
    check["sorted"] = bool((Dd[:, 1:] <= Dd[:, :-1]).all().item())
    check["ids_in_range"] = bool(((Id >= 0) & (Id < N)).all().item())

    # ---- timed: device-resident ----
    st0 = idx.stats()
    idx.set_profiling(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total = timed(step_dev, a.steps)
    clocks = sampler.stop() if rank == 0 else None
    st1 = idx.stats()
    idx.set_profiling(False)
    ms_step = ms_total / a.steps
    value = Q / (ms_step * 1e-3)
    launches = (st1["kernel_launches"] - st0["kernel_launches"]) + (a.steps if world > 1 else 0)
    kern_ms = st1["main_kernel_ms"] / max(st1["main_kernel_count"], 1)

    # ---- timed: end to end through host buffers ----
    e2e = None
    if not a.no_e2e:
        for _ in range(2):
            step_e2e()
        ms_e2e = timed(step_e2e, a.steps) / a.steps
        e2e = {"value": Q / (ms_e2e * 1e-3), "unit": "queries/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": Q * D * 4 * (world if world > 1 else 1), "d2h_bytes_per_step": Q * k * 12}

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    peaks = load_peaks()
    path = st1["last_path_name"]
    hbm_bound = path in ("stream", "tstream")
    if hbm_bound:
        esz = 4 if store == "f32" else 2
        bytes_per_launch = (hi - lo) * ((D + 63) // 64 * 64) * esz
        achieved = bytes_per_launch / (kern_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": achieved / peaks["hbm_gbs"],
                "traffic": None}
    else:
        # a step is ceil(Q / batch) launches of the same kernel over the whole shard: average work per launch
        launches_per_step = max(st1["main_kernel_count"], 1) / a.steps
        flops = 2.0 * Q * (hi - lo) * D / launches_per_step
        achieved = flops / (kern_ms * 1e-3) / 1e12
        peak = peaks["bf16_tflops_sustained"]
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                "frac_of_nominal_2250": achieved / 2250.0}
    # measured DRAM traffic of the same kernel sequence (ncu --set full, committed under profiles/), single-GPU full-size runs only
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        kname = {"tensor": "k1_search_kernel", "stream": "k2_stream_kernel", "tstream": "k2t_stream_kernel", "exact": "kx_search_kernel"}.get(path, path)
        ent = tr.get(f"{a.config}/{kname}")
        if ent and world == 1 and not (a.n or a.nq or a.k):
            roof["traffic"] = ent["bytes"]
            roof["traffic_note"] = ent["note"]
    except Exception:
        pass
    roof.update({"kernel": {"tensor": "k1_search_kernel", "stream": "k2_stream_kernel", "tstream": "k2t_stream_kernel", "exact": "kx_search_kernel"}.get(path, path),
                 "kernel_ms": kern_ms, "kernel_launches_timed": st1["main_kernel_count"], "peak_source": peaks["source"],
                 "algorithmic_work": "2*Q*(N/G)*D FLOPs per launch" if not hbm_bound else "N*ld*sizeof(T) bytes per launch"})

    cpu = None
    if world == 1 and not a.no_cpu_baseline:
        rows, sq = min(a.cpu_sample_rows, N), min(a.cpu_sample_queries, Q)
        qps, t, cores = cpu_reference_search(rows, sq, D, k, N, norm)
        cpu = {"value": qps, "unit": "queries/s", "cores": cores, "kind": "port", "seconds": t,
               "sample": f"{sq} queries x {rows} rows x {D} (of {Q} x {N}); queries/s scaled linearly in N",
               "faiss_cpu": probe_faiss_cpu(rows, sq, D, k, N, norm)}

    line = {
        "metric": "queries/sec exact top-k flat search", "value": value, "unit": "queries/s", "n_gpus": world, "steps": a.steps,
        "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": store if store != "f32" else "f32",
        "data": "synthetic",
        "config": {"workload": desc, "nq": Q, "n": N, "dim": D, "k": k, "store_dtype": store, "path": path,
                   "k_fetch": st1["last_k_fetch"], "chunks": st1["last_chunks"], "grid": st1["last_grid"],
                   "parallelism": f"corpus rows sharded x{world}" + ((", NCCL all-to-all of per-shard top-k by query block + merge kernel + all-gather of the merged blocks" if partitioned
                                                                          else ", NCCL all-gather of per-shard top-k + merge kernel") if world > 1 else ""),
                   "l2_policy": "inputs larger than L2 (corpus shard %.1f GB vs 126 MB L2); no flush needed" % ((hi - lo) * D * (4 if store == "f32" else 2) / 1e9),
                   "index_build_s": t_build},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "check": check,
    }
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    if a.impl == "reference":
        return run_reference(a)
    return run_ours(a)


if __name__ == "__main__":
    sys.exit(main())
