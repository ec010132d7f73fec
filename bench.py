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
  check   rows of the TIMED output (at N > 1: the merged output that came out of the NCCL exchange) compared with the
          oracle's exact top-k over the whole corpus, which is streamed back from the shards (nxf_reconstruct)
  cpu_baseline   the oracle's blocked numpy restatement of the reference search on a bounded sample (rank 0, N=1)
  small_batch / pool   (N=1) the memory-bound kernels on the same resident corpus / the reference's default encode batch
  c3      (--gpus 8) BASELINE config 3 on the same ranks: 100k queries x 21M x 1024, top-100, with its own oracle check

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
CORPUS_SEED = {"c1": 1001, "c2": 1002, "c2f": 1002, "c3": 1003, "c4": 1004, "c5": 1005}
QUERY_SEED = {"c1": 2001, "c2": 2002, "c2f": 2002, "c3": 2003, "c4": 2004, "c5": 2005}


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
    ap.add_argument("--dist", default="normal", choices=["normal", "clustered"], help="synthetic distribution (SURVEY 8(d): isotropic | clustered)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the small_batch / pool / c3 records")
    ap.add_argument("--check-queries", type=int, default=48, help="rows of the timed output compared with the oracle")
    ap.add_argument("--cpu-sample-rows", type=int, default=1_000_000)
    ap.add_argument("--cpu-sample-queries", type=int, default=512)
    return ap.parse_args()


# ---------------------------------------------------------------------------------------------------------------
# host threads: numpy / torch read OMP_NUM_THREADS & co. when they are first imported, and torch.distributed.run exports
# OMP_NUM_THREADS=1 to its workers -- so the count is pinned HERE, before either import (the CPU arm ran on one
# thread under torchrun and reported os.cpu_count())
# ---------------------------------------------------------------------------------------------------------------
def pin_host_threads(n: int) -> int:
    n = max(1, int(n))
    for v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
        os.environ[v] = str(n)
    return n


def host_threads_in_use() -> int:
    """What the BLAS numpy is linked to will actually use (threadpoolctl), not what was asked for."""
    try:
        import numpy  # noqa: F401
        from threadpoolctl import threadpool_info
        blas = [p["num_threads"] for p in threadpool_info() if p.get("user_api") == "blas"]
        return max(blas) if blas else 1
    except Exception:
        return int(os.environ.get("OMP_NUM_THREADS", "1"))


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


def resolve(a):
    Q, N, D, k, store, norm, desc = CONFIGS[a.config]
    if a.dim or a.store:
        D, store = a.dim or D, a.store or store
        desc += f" [override: dim {D}, store {store}]"
    if a.dist != "normal":
        desc += f" [{a.dist} data]"
    Q, N, k = a.nq or Q, a.n or N, a.k or k
    return Q, N, D, k, store, norm, desc


def make_config(a, world: int) -> dict:
    """The workload description, identical in both arms (ours / reference) for the same command line."""
    Q, N, D, k, store, norm, desc = resolve(a)
    esz = 4 if store == "f32" else 2
    shard_gb = (N / world) * D * esz / 1e9
    return {"workload": desc, "nq": Q, "n": N, "dim": D, "k": k, "store_dtype": store,
            "parallelism": f"corpus rows sharded x{world}" + (", one NCCL all-gather of packed per-shard top-k + merge kernel" if world > 1 else ""),
            "l2_policy": "inputs larger than L2 (corpus shard %.1f GB vs 126 MB L2); no flush needed" % shard_gb}


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


# ---------------------------------------------------------------------------------------------------------------
# CPU arm
# ---------------------------------------------------------------------------------------------------------------
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
    """Times the oracle port of the reference CPU search on a bounded sample and scales linearly in N to the full corpus:
    value = Q_s / (t * N / N_s).  The timed function is oracle.blocked_ip_search_thr -- the fp32 q @ block.T of NumpyIPIndex
    (utils.py:21-22) with a threshold-filtered selection, the strongest form of the port (the plain argpartition form is ~1.7x
    slower; both are asserted equal in the CPU tests).  The sgemm alone is timed too: no CPU flat search (Faiss included) beats it.
    Returns a dict."""
    import numpy as np
    import oracle
    c = oracle.make_unit_gaussian(sample_rows, d, seed=1002, normalize=normalised)
    c = oracle.round_to_bf16(c)
    q = oracle.make_unit_gaussian(sample_q, d, seed=2002, normalize=normalised)
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        oracle.blocked_ip_search_thr(c, q, k, block_rows=65536)
        if it >= warmup:
            times.append(time.perf_counter() - t0)
    t = sum(times) / len(times)
    t0 = time.perf_counter()
    for r0 in range(0, sample_rows, 65536):
        (q @ c[r0:r0 + 65536].T)
    t_gemm = time.perf_counter() - t0
    scale = n_full / float(sample_rows)
    return {"value": sample_q / (t * scale), "seconds": t, "threads": host_threads_in_use(), "host_cores": os.cpu_count() or 1,
            "sgemm_only_value": sample_q / (t_gemm * scale), "sgemm_only_seconds": t_gemm,
            "sgemm_tflops": 2.0 * sample_q * sample_rows * d / t_gemm / 1e12}


def cpu_baseline_record(a, Q, N, D, k, norm, steps=1, warmup=0):
    rows, sq = min(a.cpu_sample_rows, N), min(a.cpu_sample_queries, Q)
    r = cpu_reference_search(rows, sq, D, k, N, norm, steps=steps, warmup=warmup)
    return {"value": r["value"], "unit": "queries/s", "cores": r["threads"], "host_cores": r["host_cores"], "kind": "port",
            "seconds_per_sample": r["seconds"],
            "sample": f"{sq} queries x {rows} rows x {D} (of {Q} x {N}); queries/s scaled linearly in N",
            "sgemm_only": {"value": r["sgemm_only_value"], "unit": "queries/s", "tflops": r["sgemm_tflops"],
                           "note": "the fp32 q @ block.T of the same sample alone: an upper bound for any CPU flat search on these cores"},
            "note": "faiss is not installable offline: this is the oracle port of the reference's own NumpyIPIndex/search "
                    "(Nexus/evaluation/text_retrieval/utils.py:16-31,194-230), blocked so that it fits in RAM, with a "
                    "threshold-filtered selection; it is selection-bound (numpy), i.e. a WEAK baseline -- the roofline fraction, "
                    "not the GPU/CPU ratio, is the measure of the kernels",
            "faiss_cpu": probe_faiss_cpu(rows, sq, D, k, N, norm)}, r


def run_reference(a):
    """The reference arm: the reference's CPU implementation of the path (oracle port), all host threads, rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    Q, N, D, k, store, norm, desc = resolve(a)
    cpu, r = cpu_baseline_record(a, Q, N, D, k, norm, steps=a.steps, warmup=min(a.warmup, 1))
    line = {
        "impl": "reference", "metric": "queries/sec exact top-k flat search", "value": r["value"], "unit": "queries/s",
        "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup, "ms_per_step": r["seconds"] * 1e3, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": make_config(a, max(1, a.gpus)),
        "cpu_baseline": cpu,
        "e2e": {"value": r["value"], "unit": "queries/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


# ---------------------------------------------------------------------------------------------------------------
# synthetic data (generated on the device, never on the host)
# ---------------------------------------------------------------------------------------------------------------
def gen_rows(torch, cid_seed, rows, D, norm, device, dist="normal", centroids=None):
    g = torch.Generator(device=device)
    g.manual_seed(cid_seed)
    x = torch.randn((rows, D), generator=g, device=device, dtype=torch.float32)
    if dist == "clustered":
        # SURVEY 8(d) variant (ii): unit centroids + N(0, 0.5^2 / D) noise, renormalised
        which = torch.randint(0, centroids.shape[0], (rows,), generator=g, device=device)
        x = centroids[which] + x * (0.5 / D ** 0.5)
        return torch.nn.functional.normalize(x, dim=-1)
    if norm:
        x = torch.nn.functional.normalize(x, dim=-1)
    return x


def make_centroids(torch, D, device, n=4096):
    g = torch.Generator(device=device)
    g.manual_seed(7)
    return torch.nn.functional.normalize(torch.randn((n, D), generator=g, device=device, dtype=torch.float32), dim=-1)


def build_shard(torch, idx, lo, hi, D, store, norm, device, seed=1002, dist="normal"):
    """Synthetic corpus rows [lo, hi): N(0,1), optionally L2-normalised, rounded ONCE to the storage type
    (the stored values are the ground-truth corpus).  Generated on the device, never on the host."""
    dt = {"bf16": torch.bfloat16, "f16": torch.float16, "f32": torch.float32}[store]
    cent = make_centroids(torch, D, device) if dist == "clustered" else None
    c0, c1 = lo // GEN_CHUNK, (hi - 1) // GEN_CHUNK if hi > lo else -1
    for cid in range(c0, c1 + 1):
        x = gen_rows(torch, seed * 1_000_003 + cid, GEN_CHUNK, D, norm, device, dist, cent)
        r0 = cid * GEN_CHUNK
        a_, b_ = max(lo, r0) - r0, min(hi, r0 + GEN_CHUNK) - r0
        idx.add(x[a_:b_].to(dt))
        del x


def make_queries(torch, Q, D, norm, device, seed=2002, dist="normal"):
    cent = make_centroids(torch, D, device) if dist == "clustered" else None
    return gen_rows(torch, seed, Q, D, norm, device, dist, cent)


# ---------------------------------------------------------------------------------------------------------------
# the check of the TIMED output against the oracle
# ---------------------------------------------------------------------------------------------------------------
def pick_check_rows(Q: int, n: int, tile: int = 256):
    """Query rows spread over the first, a middle and the last query tile of the timed launch (+ the very last row)."""
    import numpy as np
    n = max(3, min(n, Q))
    per = n // 3
    starts = [0, (Q // 2) // tile * tile, max(0, (Q - 1) // tile * tile)]
    rows = []
    for s in starts:
        span = max(1, min(tile, Q - s))
        rows += [s + int(j) for j in np.linspace(0, span - 1, per).round()]
    rows.append(Q - 1)
    return sorted(set(r for r in rows if 0 <= r < Q))


def oracle_topk_over_shard(idx, q_np, k, id_offset, chunk_rows=1 << 18):
    """Exact top-k of the given queries over the rows THIS index holds, by the oracle (fp32 q @ block.T, utils.py:21-22), with the
    STORED values streamed back from the device (nxf_reconstruct) -- the storage-exact ground truth.  Global ids."""
    import numpy as np
    import oracle
    n = idx.ntotal
    best_s = np.empty((q_np.shape[0], 0), np.float32)
    best_i = np.empty((q_np.shape[0], 0), np.int64)
    for r0 in range(0, n, chunk_rows):
        rows = idx.reconstruct_n(r0, min(chunk_rows, n - r0))
        s, i = oracle.blocked_ip_search(rows, q_np, k, pad=False)
        best_s = np.concatenate([best_s, s], axis=1)
        best_i = np.concatenate([best_i, i + (r0 + id_offset)], axis=1)
        if best_s.shape[1] > k:
            order = np.lexsort((best_i, -best_s), axis=1)[:, :k]
            best_s, best_i = np.take_along_axis(best_s, order, 1), np.take_along_axis(best_i, order, 1)
    order = np.lexsort((best_i, -best_s), axis=1)
    return np.take_along_axis(best_s, order, 1), np.take_along_axis(best_i, order, 1)


def check_timed_output(torch, dist, idx_local, id_offset, q, D_timed, I_timed, k, n_check, world, rank, device):
    """Rows of the timed output vs the oracle over the whole corpus.  At world > 1 every rank reduces its own shard with the
    oracle, the per-shard lists are gathered (a checker-side collective, outside every timed region) and rank 0 merges in numpy."""
    import numpy as np
    import oracle
    t0 = time.perf_counter()
    rows = pick_check_rows(int(q.shape[0]), n_check)
    q_np = q[rows].float().cpu().numpy()
    s, i = oracle_topk_over_shard(idx_local, q_np, k, id_offset)
    if world > 1:
        kk = s.shape[1]
        pad_s = np.full((len(rows), k), -np.inf, np.float32)
        pad_i = np.full((len(rows), k), -1, np.int64)
        pad_s[:, :kk], pad_i[:, :kk] = s, i
        ts, ti = torch.from_numpy(pad_s).to(device), torch.from_numpy(pad_i).to(device)
        gs = torch.empty((world,) + tuple(ts.shape), dtype=ts.dtype, device=device)
        gi = torch.empty((world,) + tuple(ti.shape), dtype=ti.dtype, device=device)
        dist.all_gather_into_tensor(gs, ts)
        dist.all_gather_into_tensor(gi, ti)
        if rank != 0:
            return None
        cs = gs.permute(1, 0, 2).reshape(len(rows), -1).cpu().numpy()
        ci = gi.permute(1, 0, 2).reshape(len(rows), -1).cpu().numpy()
        order = np.lexsort((ci, -cs), axis=1)[:, :k]
        s, i = np.take_along_axis(cs, order, 1), np.take_along_axis(ci, order, 1)
        s = np.where(i >= 0, s, -oracle.flat_search.FLT_MAX).astype(np.float32)
    else:
        s, i = oracle.faiss_pad(s, i, k)
    got_s = D_timed[rows].float().cpu().numpy()
    got_i = I_timed[rows].cpu().numpy()
    rep = oracle.check_topk(s, i, got_s, got_i, None, None, rtol=1e-5, atol=1e-6)
    return {"source": "timed_output", "oracle_queries": len(rows), "oracle_rows": rows[:3] + ["..."] + rows[-3:],
            "oracle_recall": rep.recall, "oracle_bad_ids": rep.bad_ids, "oracle_tie_swaps": rep.tie_swaps,
            "oracle_max_rel_err": rep.max_score_rel_err, "oracle_exact_id_rows": rep.exact_id_rows, "oracle_ok": bool(rep.ok),
            "oracle_rtol": 1e-5, "oracle_seconds": time.perf_counter() - t0,
            "oracle_how": "oracle.blocked_ip_search (fp32 q @ block.T, utils.py:21-22) over the stored rows streamed back with nxf_reconstruct"
                          + ("; per-shard lists merged in numpy on rank 0" if world > 1 else ""),
            "oracle_messages": rep.messages[:4]}


# ---------------------------------------------------------------------------------------------------------------
# our arm
# ---------------------------------------------------------------------------------------------------------------
def timed_steps(torch, dist, fn, steps, world, device):
    def barrier():
        if world > 1:
            dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    torch.cuda.synchronize()
    e0.record()
    out = None
    for _ in range(steps):
        out = fn()
    e1.record()
    torch.cuda.synchronize()
    barrier()
    ms = torch.tensor([e0.elapsed_time(e1)], device=device, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item()), out


def small_batch_records(torch, idx, q, k, N, D, store, peaks, nqs=(1, 32, 128), iters=10):
    """The memory-bound path on the resident corpus, the way the reference CALLS search (serve nq = 1, eval chunk 32, recsys 128):
    kernel time from the library's CUDA events, achieved GB/s = N*ld*sizeof(T) / kernel time."""
    out = []
    ld = (D + 63) // 64 * 64
    for nq in nqs:
        qq = q[:nq].contiguous()
        for _ in range(3):
            idx.search(qq, k)
        torch.cuda.synchronize()
        idx.set_profiling(True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            idx.search(qq, k)
        e1.record()
        torch.cuda.synchronize()
        st = idx.stats()
        idx.set_profiling(False)
        kms = st["main_kernel_ms"] / max(1, st["main_kernel_count"])
        call = e0.elapsed_time(e1) / iters
        # bytes the kernel streams per corpus element: the tensor paths read 16-bit rows (fp32 storage: its bf16 plane), K2 / KX the rows as stored
        esz = 2 if (store != "f32" or st["last_path_name"] in ("tstream", "tensor")) else 4
        gbs = N * ld * esz / (kms * 1e-3) / 1e9
        out.append({"nq": nq, "k": k, "path": st["last_path_name"], "kernel_ms": kms, "call_ms": call, "host_and_merge_ms": call - kms,
                    "streamed_bytes_per_element": esz, "GBps": gbs, "frac": gbs / peaks["hbm_gbs"], "queries_per_s": nq / (call * 1e-3)})
    return out


def pool_record(torch, peaks, device, B=256, L=512, H=768, iters=20):
    """K0 on the reference's default encode batch (base.py:59,355-379: 256 x 512 x 768 fp16, mean pooling + normalise):
    algorithmic bytes = unmasked positions * H * 2 + B*L*8 + B*H*4 over the CUDA-event time per call."""
    from nexus_b200.pooling import pool_and_normalize
    g = torch.Generator(device=device)
    g.manual_seed(7)
    n_buf = 3          # 3 x 201 MB > L2: consecutive calls do not find their input in L2
    hs = [torch.randn((B, L, H), generator=g, device=device, dtype=torch.float32).to(torch.float16) for _ in range(n_buf)]
    lens = torch.randint(L // 2, L + 1, (B,), generator=g, device=device)
    mask = (torch.arange(L, device=device)[None, :] < lens[:, None]).long()
    kept = int(mask.sum().item())
    alg = kept * H * 2 + B * L * 8 + B * H * 4
    for i in range(3):
        pool_and_normalize(hs[i % n_buf], mask, "mean")
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        pool_and_normalize(hs[i % n_buf], mask, "mean")
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / iters * 1e3
    gbs = alg / (us * 1e-6) / 1e9
    return {"shape": f"{B} x {L} x {H} f16, mean + L2-normalise, right-padded masks ({kept} of {B * L} positions kept)", "us": us,
            "algorithmic_bytes": alg, "GBps": gbs, "frac": gbs / peaks["hbm_gbs"], "l2_policy": "3 input buffers of 201 MB rotate (> 126 MB L2)"}


def run_c3(torch, dist, world, rank, local_rank, device, peaks, n_check):
    """BASELINE config 3 on the same ranks: 100k x 21M x 1024 bf16, top-100, rows sharded, 3 timed steps + oracle check."""
    from nexus_b200 import DistributedFlatIndex, shard_bounds
    Q, N, D, k, store, norm, desc = CONFIGS["c3"]
    lo, hi = shard_bounds(N, world)[rank]
    didx = DistributedFlatIndex(D, "ip", store, device=local_rank)
    didx.local.reserve(hi - lo)
    didx.local.set_id_offset(lo)
    build_shard(torch, didx.local, lo, hi, D, store, norm, device, seed=CORPUS_SEED["c3"])
    didx._ntotal_global = N
    didx._offsets = [b[0] for b in shard_bounds(N, world)]
    q = make_queries(torch, Q, D, norm, device, seed=QUERY_SEED["c3"])
    for _ in range(2):
        didx.search(q, k)
    didx.local.set_profiling(True)
    ms, out = timed_steps(torch, dist, lambda: didx.search(q, k), 3, world, device)
    st = didx.local.stats()
    didx.local.set_profiling(False)
    chk = check_timed_output(torch, dist, didx.local, lo, q, out[0], out[1], k, n_check, world, rank, device)
    if rank != 0:
        return None
    ms_step = ms / 3
    kern_ms = st["main_kernel_ms"] / 3          # summed over the launches of a step (a 100k batch is several launches)
    tf = 2.0 * Q * (hi - lo) * D / (kern_ms * 1e-3) / 1e12
    return {"workload": desc, "nq": Q, "n": N, "dim": D, "k": k, "steps": 3, "ms_per_step": ms_step, "value": Q / (ms_step * 1e-3),
            "unit": "queries/s", "kernel_ms_per_step": kern_ms, "per_gpu_tflops": tf, "frac": tf / peaks["bf16_tflops_sustained"],
            "whole_job_tflops_per_gpu": 2.0 * Q * (hi - lo) * D / (ms_step * 1e-3) / 1e12, "check": chk}


def run_ours(a):
    import numpy as np  # noqa: F401
    import torch
    import torch.distributed as dist
    from nexus_b200 import DistributedFlatIndex, FlatIndex, shard_bounds
    from nexus_b200 import _lib as L

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

    Q, N, D, k, store, norm, desc = resolve(a)
    lo, hi = shard_bounds(N, world)[rank]
    if world > 1:
        didx = DistributedFlatIndex(D, "ip", store, device=local_rank)
        idx = didx.local
        didx._ntotal_global = N
        didx._offsets = [b[0] for b in shard_bounds(N, world)]
    else:
        didx = None
        idx = FlatIndex(D, "ip", store, device=local_rank)
    idx.reserve(hi - lo)
    idx.set_id_offset(lo)
    t_build = time.perf_counter()
    build_shard(torch, idx, lo, hi, D, store, norm, device, seed=CORPUS_SEED[a.config], dist=a.dist)
    torch.cuda.synchronize()
    t_build = time.perf_counter() - t_build

    q = make_queries(torch, Q, D, norm, device, seed=QUERY_SEED[a.config], dist=a.dist)
    q_host = q.cpu().pin_memory()
    out_d = torch.empty((Q, k), dtype=torch.float32).pin_memory()
    out_i = torch.empty((Q, k), dtype=torch.int64).pin_memory()
    flags = a.flags
    idx.search_flags = flags
    exchange = os.environ.get("NXF_BENCH_EXCHANGE", "auto")      # A/B of the two exchanges: gather | partition
    partitioned = world > 2 and exchange != "gather" and (exchange == "partition" or Q * k * 8 * world >= (256 << 20))

    def step_dev():
        if world == 1:
            return idx.search(q, k)
        return didx.search(q, k, exchange="partition" if partitioned else "gather")

    def step_e2e():
        if world == 1:
            return idx.search(q_host, k, out=(out_d, out_i))     # host buffers through the C ABI
        qd = q_host.to(device, non_blocking=True)
        D_, I_ = didx.search(qd, k, exchange="partition" if partitioned else "gather")
        if rank == 0:
            out_d.copy_(D_, non_blocking=True)
            out_i.copy_(I_, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        return out_d, out_i

    # ---- warm-up (>= 3) ----
    W = max(a.warmup, 3)
    for _ in range(W):
        step_dev()
    torch.cuda.synchronize()

    # ---- timed: device-resident ----
    st0 = idx.stats()
    idx.set_profiling(True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    ms_total, (Dd, Id) = timed_steps(torch, dist, step_dev, a.steps, world, device)
    clocks = sampler.stop() if rank == 0 else None
    st1 = idx.stats()
    idx.set_profiling(False)
    ms_step = ms_total / a.steps
    value = Q / (ms_step * 1e-3)
    launches = (st1["kernel_launches"] - st0["kernel_launches"])      # the library counts its own launches (K3 merge: + 1 per step)
    if world > 1:
        launches += a.steps
    kern_ms = st1["main_kernel_ms"] / max(st1["main_kernel_count"], 1)

    # ---- the TIMED output against the oracle (rows of the last timed step; at N > 1 the merged output of the exchange) ----
    check = check_timed_output(torch, dist, idx, lo, q, Dd, Id, k, a.check_queries, world, rank, device) or {}
    check["sorted"] = bool((Dd[:, 1:] <= Dd[:, :-1]).all().item())
    check["ids_in_range"] = bool(((Id >= 0) & (Id < N)).all().item())
    if world == 1:
        # second key: the tensor path against the on-device exact fp32 kernel (KX) on 64 queries over the full corpus
        Dl, Il = Dd[:64], Id[:64]
        De, Ie = idx.search(q[:64], k, flags=L.SEARCH_FORCE_EXACT)
        inter = sum(len(set(Il[r].tolist()) & set(Ie[r].tolist())) for r in range(Il.shape[0]))
        check["recall_vs_exact_fp32_kernel_64q"] = inter / float(Il.numel())
        check["max_abs_score_diff_top1"] = float((Dl[:, 0] - De[:, 0]).abs().max().item())

    # ---- timed: end to end through host buffers ----
    e2e = None
    if not a.no_e2e:
        for _ in range(2):
            step_e2e()
        ms_e2e = timed_steps(torch, dist, step_e2e, a.steps, world, device)[0] / a.steps
        e2e = {"value": Q / (ms_e2e * 1e-3), "unit": "queries/s", "ms_per_step": ms_e2e,
               "h2d_bytes_per_step": Q * D * 4 * (world if world > 1 else 1), "d2h_bytes_per_step": Q * k * 12}

    peaks = load_peaks()
    small = pool = c3 = None
    if not a.no_extras and not (a.n or a.nq or a.k or a.dim or a.store or a.flags):
        if world == 1:
            small = small_batch_records(torch, idx, q, k, N, D, store, peaks)
            pool = pool_record(torch, peaks, device)
        # NXF_BENCH_C3=Q,N forces the record at any rank count with a reduced shape (a dry run of this code path on 2 GPUs)
        c3_env = os.environ.get("NXF_BENCH_C3", "")
        if (world == 8 or c3_env) and world > 1 and a.config == "c2":
            if c3_env:
                cq, cn = (int(x) for x in c3_env.split(","))
                _, _, cD, ck, cstore, cnorm, cdesc = CONFIGS["c3"]
                CONFIGS["c3"] = (cq, cn, cD, ck, cstore, cnorm, cdesc + f" [reduced: {cq} x {cn}]")
            del idx, didx
            torch.cuda.empty_cache()
            c3 = run_c3(torch, dist, world, rank, local_rank, device, peaks, a.check_queries)

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

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
                "frac_of_nominal_2250": achieved / 2250.0, "frac_of_burst_peak": achieved / peaks["bf16_tflops"],
                "peak_note": "peak = the SUSTAINED cuBLAS bf16 figure of MEASURED_PEAKS.json (a kernel timed inside a long, power-capped step); "
                             "frac_of_burst_peak is against the best-of-10 figure of the same file"}
    # DRAM traffic of the same kernel sequence: NOT measured in this run -- copied from the committed ncu --set full capture
    try:
        tr = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        kname = {"tensor": "k1_search_kernel", "stream": "k2_stream_kernel", "tstream": "k2t_stream_kernel", "exact": "kx_search_kernel"}.get(path, path)
        ent = tr.get(f"{a.config}/{kname}")
        if ent and world == 1 and not (a.n or a.nq or a.k):
            roof["traffic"] = ent["bytes"]
            roof["traffic_note"] = ent["note"]
            roof["traffic_source"] = ent.get("source", "profiles/traffic.json (ncu --set full capture of an earlier run of this command; not live)")
    except Exception:
        pass
    roof.update({"kernel": {"tensor": "k1_search_kernel", "stream": "k2_stream_kernel", "tstream": "k2t_stream_kernel", "exact": "kx_search_kernel"}.get(path, path),
                 "kernel_ms": kern_ms, "kernel_launches_timed": st1["main_kernel_count"], "peak_source": peaks["source"],
                 "algorithmic_work": "2*Q*(N/G)*D FLOPs per launch" if not hbm_bound else "N*ld*sizeof(T) bytes per launch"})

    cpu = None
    if world == 1 and not a.no_cpu_baseline:
        cpu, _ = cpu_baseline_record(a, Q, N, D, k, norm)

    line = {
        "metric": "queries/sec exact top-k flat search", "value": value, "unit": "queries/s", "n_gpus": world, "steps": a.steps,
        "warmup": W, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": store,
        "data": "synthetic",
        "config": make_config(a, world),
        "plan": {"path": path, "k_fetch": st1["last_k_fetch"], "chunks": st1["last_chunks"], "grid": st1["last_grid"], "index_build_s": t_build,
                 "exchange": ("partition" if partitioned else "gather") if world > 1 else None, "host_threads": host_threads_in_use(),
                 # certified search over the timed steps (rank 0's shard): queries checked / searched again by the segmented pass / by the exact kernels
                 "certificate": {"checked": st1["cert_checked"] - st0["cert_checked"], "retry_segmented": st1["cert_retry_l2"] - st0["cert_retry_l2"],
                                 "retry_exact": st1["cert_retry_l3"] - st0["cert_retry_l3"]}},
        "clocks": clocks, "e2e": e2e, "gpu_launches": launches, "roofline": roof, "cpu_baseline": cpu, "check": check,
    }
    if small is not None:
        line["small_batch"] = small
    if pool is not None:
        line["pool"] = pool
    if c3 is not None:
        line["c3"] = c3
    print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    a = parse_args()
    cores = os.cpu_count() or 1
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("WORLD_SIZE", "1")))
    if a.impl == "reference":
        pin_host_threads(cores)                         # rank 0 alone works: all host threads, whatever the launcher exported
        return run_reference(a)
    pin_host_threads(max(1, cores // max(1, local_world)))
    return run_ours(a)


if __name__ == "__main__":
    sys.exit(main())
