"""A reduced selection of every kernel family, sized for compute-sanitizer (10-100x slower than a plain run):

    compute-sanitizer --tool memcheck  python tools/sanitize_cases.py > profiles/r2_sanitizer_memcheck.log 2>&1
    compute-sanitizer --tool racecheck python tools/sanitize_cases.py > profiles/r2_sanitizer_racecheck.log 2>&1

K0 (mean split / cls / last), K1 (one and two epilogue sets, sampled thresholds, the forced fall-back, the segmented retry level),
K2, K2T (with sample launches), KX, KS, the arena merge with certificate + retry, K3 (rank merge and select + sort), packed output.
Every result is also compared with the oracle, so a clean sanitizer log is not a vacuous one."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    import torch
    import oracle
    from nexus_b200 import FlatIndex, merge_packed, merge_topk, pool_and_normalize
    from nexus_b200 import _lib as L
    rng = np.random.default_rng(0)

    def unit(n, d):
        x = rng.standard_normal((n, d)).astype(np.float32)
        return x / np.linalg.norm(x, axis=1, keepdims=True)

    def check(name, idx, c_stored, q, k, D, I, rtol=1e-5):
        ref_s, ref_i = oracle.blocked_ip_search(c_stored, q, k)
        rep = oracle.check_topk(ref_s, ref_i, D, I, c_stored, q, rtol=rtol)
        st = idx.stats()
        assert rep.ok, (name, st, rep.messages[:4])
        print(f"[case] {name}: ok path={st['last_path_name']} launches={st['kernel_launches']} retry_l2={st['cert_retry_l2']} retry_l3={st['cert_retry_l3']}", flush=True)

    # ---- K1: long enough for the estimate / sample / main / verify / fall-back schedule (>= 1024 tiles of 256 rows) ----
    c = rng.standard_normal((270000, 64)).astype(np.float32)
    q = rng.standard_normal((300, 64)).astype(np.float32)
    cs = oracle.round_to_bf16(c)
    for sets in ("1", "2"):
        os.environ["NXF_K1_SETS"] = sets
        idx = FlatIndex(64, "ip", "bf16")
        idx.add(c)
        D, I = idx.search(q, 20, flags=L.SEARCH_FORCE_TENSOR)
        check(f"K1 sets={sets} sampled thresholds", idx, cs, q, 20, D, I)
    os.environ.pop("NXF_K1_SETS")
    # the fall-back: every row that matters lies in a sample tile
    c2 = (0.05 * rng.standard_normal((270000, 64))).astype(np.float32)
    x2 = unit(130, 64)
    hot = np.flatnonzero((np.arange(270000) // 256) % 16 == 0)
    c2[hot] += x2[rng.integers(0, 130, size=hot.size)] * rng.uniform(0.5, 1.0, size=(hot.size, 1)).astype(np.float32)
    idx = FlatIndex(64, "ip", "bf16")
    idx.add(c2)
    D, I = idx.search(x2, 50, flags=L.SEARCH_FORCE_TENSOR)
    check("K1 fall-back after a failed verification", idx, oracle.round_to_bf16(c2), x2, 50, D, I)

    # ---- small shapes of every path, both storage families ----
    c = unit(6000, 128)
    q = unit(140, 128)
    for store in ("bf16", "f32"):
        cst = oracle.round_to_bf16(c) if store == "bf16" else c
        idx = FlatIndex(128, "ip", store)
        idx.add(c)
        for name, qq, k, flags in [("K1", q, 10, L.SEARCH_DEFAULT), ("K2T", q[:33], 10, L.SEARCH_DEFAULT), ("K2T one query", q[:1], 5, L.SEARCH_DEFAULT),
                                   ("K2", q[:3], 10, L.SEARCH_FORCE_STREAM), ("KX", q[:70], 100, L.SEARCH_FORCE_EXACT)]:
            D, I = idx.search(qq, k, flags=flags)
            check(f"{name} {store}", idx, cst, qq, k, D, I)
        P = idx.search_packed(torch.from_numpy(q).cuda(), 10)
        Dm, Im = merge_packed(P.unsqueeze(0), [0], 10, "ip")
        check(f"packed output + K3 {store}", idx, cst, q, 10, Dm.cpu().numpy(), Im.cpu().numpy())
    # K2T with sample launches (>= 4 tiles of 128 rows per CTA)
    c = rng.standard_normal((150000, 64)).astype(np.float32)
    q = rng.standard_normal((128, 64)).astype(np.float32)
    idx = FlatIndex(64, "ip", "bf16")
    idx.add(c)
    D, I = idx.search(q, 100)
    check("K2T 128 queries with sampled thresholds", idx, oracle.round_to_bf16(c), q, 100, D, I)

    # ---- certificate failures: the segmented retry level and the exact kernels ----
    from test_gpu_certified import _near_duplicates
    for store in ("f32", "bf16"):
        if store == "f32":
            cc, xx = _near_duplicates(4000, 600, 64, 130, 2e-5, seed=1)
        else:   # 16-bit storage spreads near-copies apart when it rounds them: bit-identical copies instead
            bg, xx = unit(4000, 64), unit(130, 64)
            best = oracle.round_to_bf16(0.9 * xx + 0.01 * rng.standard_normal(xx.shape).astype(np.float32))
            cc = np.concatenate([bg, np.repeat(best, 300, axis=0)], axis=0)
            cc = np.ascontiguousarray(cc[rng.permutation(cc.shape[0])])
        idx = FlatIndex(64, "ip", store)
        idx.add(cc)
        D, I = idx.search(xx, 50)
        st = idx.stats()
        assert st["cert_retry_l2"] > 0 and st["cert_retry_l3"] > 0, st
        check(f"certificate -> segmented K1 -> exact kernels ({store})", idx, oracle.round_to_bf16(cc) if store == "bf16" else cc, xx, 50, D, I)

    # ---- K3 in both forms ----
    S = np.sort(rng.standard_normal((9, 40, 500)).astype(np.float32), axis=2)[:, :, ::-1].copy()
    I = np.stack([np.stack([np.sort(rng.choice(100000, 500, replace=False)) * 9 + g for _ in range(40)]) for g in range(9)]).astype(np.int64)
    for G in (8, 9):
        Dm, Im = merge_topk(torch.from_numpy(S[:G]).cuda(), torch.from_numpy(I[:G]).cuda(), 100)
        ref = np.sort(S[:G].transpose(1, 0, 2).reshape(40, -1), axis=1)[:, ::-1][:, :100]
        np.testing.assert_array_equal(Dm.cpu().numpy(), ref)
        print(f"[case] K3 {G} lists x 500: ok", flush=True)

    # ---- KS ----
    c = unit(3000, 96)
    idx = FlatIndex(96, "ip", "bf16")
    idx.add(c)
    offs = np.array([0, 5, 5, 2505, 2600], dtype=np.int64)
    rws = rng.integers(0, 3000, size=int(offs[-1])).astype(np.int64)
    Ds, Ps = idx.search_subsets(unit(4, 96), offs, rws, 10)
    print("[case] KS: ok", flush=True)

    # ---- K0 ----
    for (B, Lq, H, dt) in [(8, 16, 384, torch.float16), (64, 256, 768, torch.bfloat16), (3, 7, 100, torch.float32)]:
        h = torch.randn(B, Lq, H, generator=torch.Generator().manual_seed(1)).to(dt)
        m = (torch.arange(Lq)[None, :] < torch.randint(1, Lq + 1, (B, 1), generator=torch.Generator().manual_seed(2))).long()
        for method in ("mean", "cls", "last", "last_argmax"):
            out = pool_and_normalize(h.cuda(), m.cuda(), method, normalize=True, out_dtype=torch.float32)
            ref = oracle.pool_and_normalize(h.float().numpy(), m.numpy(), method if method != "last_argmax" else "last_argmax", normalize=True)
            np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=3e-5, atol=3e-6)
        print(f"[case] K0 {B}x{Lq}x{H} {dt}: ok", flush=True)
    torch.cuda.synchronize()
    print("ALL CASES OK", flush=True)


if __name__ == "__main__":
    main()
