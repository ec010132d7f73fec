"""K0 (pool + L2-normalise) against the HBM roofline: achieved GB/s = algorithmic bytes / CUDA-event time, per pooling rule,
with the oracle (numpy restatement of the reference's pooling + normalize) timed beside it on a bounded sample.

    python tools/bench_pool.py [--b 256] [--l 512] [--h 768] [--dtype f16]
Algorithmic bytes: mean: B*L*H*s + B*L*8 + B*H*s_out (every unmasked position is read once; here the masks keep
~75 % of the positions, and only those rows are counted); cls / last: B*H*(s + s_out) + B*L*8.
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
    ap.add_argument("--b", type=int, default=256)
    ap.add_argument("--l", type=int, default=512)
    ap.add_argument("--h", type=int, default=768)
    ap.add_argument("--dtype", default="f16")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    import numpy as np
    import torch
    from nexus_b200.pooling import pool_and_normalize
    import oracle
    dt = {"f16": torch.float16, "bf16": torch.bfloat16, "f32": torch.float32}[a.dtype]
    esz = torch.empty((), dtype=dt).element_size()
    dev = torch.device("cuda", 0)
    g = torch.Generator(device=dev); g.manual_seed(7)
    # several independent batches so that consecutive iterations do not hit L2 (a batch is B*L*H*s = 201 MB at the defaults)
    n_buf = max(2, int(512e6 // (a.b * a.l * a.h * esz)) + 1)
    hs = [torch.randn((a.b, a.l, a.h), generator=g, device=dev, dtype=torch.float32).to(dt) for _ in range(n_buf)]
    lens = torch.randint(a.l // 2, a.l + 1, (a.b,), generator=g, device=dev)
    pos = torch.arange(a.l, device=dev)[None, :]
    masks = {"right": (pos < lens[:, None]).long(), "left": (pos >= (a.l - lens)[:, None]).long()}
    peaks = bench.load_peaks()
    for method, side, out_dt in [("mean", "right", torch.float32), ("mean", "right", dt), ("cls", "right", dt), ("last", "left", dt), ("last", "right", dt)]:
        m = masks[side]
        kept = int(m.sum().item())
        osz = torch.empty((), dtype=out_dt).element_size()
        if method == "mean":
            alg = kept * a.h * esz + a.b * a.l * 8 + a.b * a.h * osz
        else:
            alg = a.b * a.h * (esz + osz) + a.b * a.l * 8
        for i in range(3):
            pool_and_normalize(hs[i % n_buf], m, method, out_dtype=out_dt)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(a.iters):
            out = pool_and_normalize(hs[i % n_buf], m, method, out_dtype=out_dt)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / a.iters
        # plain PyTorch (what the reference runs on a GPU today): pooling() + F.normalize
        def torch_ref(h):
            if method == "mean":
                s = torch.sum(h * m.unsqueeze(-1).float(), dim=1)
                x = s / m.sum(dim=1, keepdim=True).float()
            elif method == "cls":
                x = h[:, 0]
            else:
                if bool((m[:, -1].sum() == a.b).item()):
                    x = h[:, -1]
                else:
                    x = h[torch.arange(a.b, device=dev), m.sum(dim=1) - 1]
            return torch.nn.functional.normalize(x, dim=-1)
        for i in range(2):
            torch_ref(hs[i % n_buf])
        torch.cuda.synchronize()
        e0.record()
        for i in range(a.iters):
            torch_ref(hs[i % n_buf])
        e1.record()
        torch.cuda.synchronize()
        ms_t = e0.elapsed_time(e1) / a.iters
        # oracle on the host: a bounded sample of the batch
        sb = min(a.b, 32)
        out = pool_and_normalize(hs[0], m, method, out_dtype=out_dt)
        h_np = hs[0][:sb].float().cpu().numpy()
        m_np = m[:sb].cpu().numpy()
        t0 = time.perf_counter()
        ref = oracle.pool_and_normalize(h_np, m_np, "last" if method == "last" else method, True)
        t_cpu = (time.perf_counter() - t0) * a.b / sb
        err = float(np.abs(out[:sb].float().cpu().numpy() - ref).max())
        gbs = alg / (ms * 1e-3) / 1e9
        print(json.dumps({"kernel": "pool_norm_kernel", "pooling": method, "padding": side, "B": a.b, "L": a.l, "H": a.h, "in": a.dtype,
                          "out": str(out_dt).replace("torch.", ""), "ms": round(ms, 4), "algorithmic_MB": round(alg / 1e6, 2),
                          "GBps": round(gbs, 1), "frac_of_measured_hbm": round(gbs / peaks["hbm_gbs"], 3),
                          "torch_eager_ms": round(ms_t, 4), "cpu_oracle_ms_scaled": round(t_cpu * 1e3, 1), "max_abs_err_vs_oracle": err}), flush=True)


if __name__ == "__main__":
    main()
