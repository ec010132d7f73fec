"""First-contact GPU diagnostics: each kernel family on small shapes, every case in its own subprocess
with a timeout (a hung mbarrier protocol must not take the whole gpurun call down).

    python tools/gpu_diag.py [--only exact,stream,tensor1,tensor2,pool,merge] [--timeout 120]
Writes gpurun_out/diag.log.
"""
import argparse
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CASES = {
    "exact": [("f32", "exact", 3000, 64, 37, 10), ("bf16", "exact", 9000, 768, 257, 100), ("f32", "exact", 4096, 96, 20, 500)],
    "stream": [("f32", "stream", 3000, 64, 1, 10), ("bf16", "stream", 20000, 384, 4, 100), ("f16", "stream", 9000, 768, 3, 20),
               ("f32", "stream", 7001, 1536, 2, 10)],
    "tensor1": [("bf16", "tensor_raw", 600, 64, 100, 10), ("bf16", "tensor_raw", 3000, 64, 37, 10), ("bf16", "tensor_raw", 9000, 768, 257, 100),
                ("f16", "tensor_raw", 5000, 128, 130, 100), ("bf16", "tensor", 20000, 384, 300, 10), ("bf16", "tensor", 4096, 96, 20, 500)],
}
CASES["tensor2"] = CASES["tensor1"]


def child(spec):
    import numpy as np
    import oracle
    from _common import gen_inputs
    from _gpu_common import FLAGS, stored
    from nexus_b200 import FlatIndex
    store, path, n, d, q, k = spec
    c, x = gen_inputs(n, d, q, seed=n + d)
    idx = FlatIndex(d, "ip", store)
    idx.add(c)
    t0 = time.time()
    D, I = idx.search(x, k, flags=FLAGS[path])
    dt = time.time() - t0
    cs = stored(c, store)
    xs = stored(x, store) if path == "tensor_raw" else x
    ref_s, ref_i = oracle.blocked_ip_search(cs, xs, k)
    rtol, atol = (2e-4, 2e-5) if path == "tensor_raw" else (1e-5, 1e-6)
    rep = oracle.check_topk(ref_s, ref_i, D, I, cs, xs, rtol=rtol, atol=atol)
    out = dict(spec=spec, ok=bool(rep.ok), recall=rep.recall, exact_rows=rep.exact_id_rows, n_rows=rep.n_rows, bad=rep.bad_ids,
               ties=rep.tie_swaps, max_rel=rep.max_score_rel_err, unsorted=rep.unsorted_rows, ms=dt * 1e3, stats=idx.stats(),
               msgs=rep.messages[:4])
    if not rep.ok:
        out["got_s0"] = D[0, :8].tolist()
        out["ref_s0"] = ref_s[0, :8].tolist()
        out["got_i0"] = I[0, :8].tolist()
        out["ref_i0"] = ref_i[0, :8].tolist()
    print("RESULT " + json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="exact,stream,tensor1,tensor2")
    ap.add_argument("--timeout", type=int, default=120)
    ap.add_argument("--child", default=None)
    a = ap.parse_args()
    if a.child:
        child(json.loads(a.child))
        return
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    log = open(os.path.join(ROOT, "gpurun_out", "diag.log"), "a")
    for group in a.only.split(","):
        env = dict(os.environ)
        env["NXF_K1_CTA_GROUP"] = "2" if group == "tensor2" else "1"
        for spec in CASES[group]:
            t0 = time.time()
            try:
                r = subprocess.run([sys.executable, __file__, "--child", json.dumps(spec)], capture_output=True, text=True,
                                   timeout=a.timeout, env=env)
                lines = [l for l in r.stdout.splitlines() if l.startswith("RESULT ")]
                msg = lines[-1] if lines else f"NO RESULT rc={r.returncode} stderr={r.stderr[-1500:]}"
            except subprocess.TimeoutExpired:
                msg = f"TIMEOUT after {a.timeout}s"
            line = f"[{group}] {spec} ({time.time() - t0:.1f}s): {msg}"
            print(line, flush=True)
            log.write(line + "\n")
            log.flush()


if __name__ == "__main__":
    main()
