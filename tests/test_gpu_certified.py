"""GPU parity of the CERTIFIED search (include/nexus_flat.h "Exactness"): candidate pass on 16-bit operands with
k' > k, fp32 re-score, certificate s_k >= tau' + eps, widen-and-retry through the segmented tensor-core pass and the exact fp32 kernels.
Every result is compared with the oracle (fp32 arithmetic on the stored values, utils.py:16-31); the counters of nxf_stats say which
levels ran."""
import numpy as np
import pytest

import oracle
from _common import gen_inputs
from _gpu_common import FLAGS, stored

pytestmark = pytest.mark.gpu


def _check(idx, c, x, k, store, D, I, rtol=1e-5):
    cs = stored(c, store)
    ref_s, ref_i = oracle.blocked_ip_search(cs, x.astype(np.float32), k)
    ref_s, ref_i = ref_s[:, :D.shape[1]], ref_i[:, :D.shape[1]]          # k > n: the caller passes the min(k, n) real columns
    rep = oracle.check_topk(ref_s, ref_i, D, I, cs, x.astype(np.float32), rtol=rtol)
    assert rep.ok, (idx.stats(), rep.messages[:6], rep)
    return rep


def _near_duplicates(n_bg, n_dup, d, q, spread, seed, renorm=True):
    """A background of random unit rows plus, for every query, a cloud of rows whose scores with it lie within `spread` of each other:
    the gap between ranks k and k' inside the cloud is what the certificate tests."""
    rng = np.random.default_rng(seed)
    bg = rng.standard_normal((n_bg, d)).astype(np.float32)
    bg /= np.linalg.norm(bg, axis=1, keepdims=True)
    x = rng.standard_normal((q, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    clouds = []
    for i in range(q):
        # rows = (1 - delta) * x_i + small orthogonal noise; delta uniform in [0, spread]
        delta = rng.uniform(0, spread, size=(n_dup, 1)).astype(np.float32)
        noise = rng.standard_normal((n_dup, d)).astype(np.float32) * np.float32(1e-3)
        noise -= (noise @ x[i])[:, None] * x[i][None, :]
        clouds.append((1 - delta) * x[i][None, :] + noise)
    c = np.concatenate([bg] + clouds, axis=0)
    perm = rng.permutation(c.shape[0])
    return np.ascontiguousarray(c[perm]), x


@pytest.mark.parametrize("store,case", [("f32", (60000, 128, 300, 100)), ("f32", (20000, 768, 200, 10)), ("bf16", (60000, 128, 300, 100)),
                                        ("f16", (30000, 64, 150, 50)), ("f32", (500, 64, 140, 100)), ("f32", (90, 64, 130, 100))])
def test_isotropic_inputs_pass_at_level_one(store, case):
    """Random unit vectors: the gap between ranks k and k' is far above the bound -- every query is certified by the candidate pass."""
    from nexus_b200 import FlatIndex
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=900 + d)
    idx = FlatIndex(d, "ip", store)
    idx.add(c)
    D, I = idx.search(x, k)
    st = idx.stats()
    assert st["last_path_name"] == "tensor" and st["cert_checked"] == q, st
    assert st["cert_retry_l3"] == 0, st
    if n >= 20000:
        assert st["cert_retry_l2"] == 0, st
        want = max(4 * k, k + 64) if store == "f32" else max(2 * k, k + 16)       # k' of the candidate pass, rounded up to 8
        assert st["last_k_fetch"] == (want + 7) // 8 * 8, st
    _check(idx, c, x, k, store, D[:, :min(k, n)], I[:, :min(k, n)])


@pytest.mark.parametrize("store", ["f32", "bf16"])
@pytest.mark.parametrize("nq", [150, 40, 1])
def test_level_two_resolves_moderately_close_scores(store, nq):
    """Clouds of 2 000 rows per query with scores spread over 0.02: ranks 100..400 lie 0.003 apart (below the level-1 bound 0.0039 /
    0.002 for 16-bit storage is borderline), ranks 100..128 2.8e-4 apart (above the level-2 bound): the segmented pass settles it."""
    from nexus_b200 import FlatIndex
    spread = 0.02 if store == "f32" else 0.006
    c, x = _near_duplicates(20000, 2000, 64, nq, spread, seed=5)
    idx = FlatIndex(64, "ip", store)
    idx.add(c)
    D, I = idx.search(x, 100)
    st = idx.stats()
    assert st["cert_checked"] >= nq and st["cert_retry_l2"] > 0, st
    assert st["last_path_name"] == ("tensor" if nq > 128 else "tstream"), st          # the stats describe the primary level
    _check(idx, c, x, 100, store, D, I)


@pytest.mark.parametrize("store", ["f32", "bf16"])
@pytest.mark.parametrize("mode,nq", [("numpy", 140), ("cuda", 140), ("packed", 140), ("numpy", 5), ("cuda", 3)])
def test_near_duplicates_force_the_exact_level(store, mode, nq):
    """Rows that no bound separates at rank k: 1 000 near-copies of every query's best row with scores within 2e-5 of each other
    (fp32 storage), 400 bit-identical copies (16-bit storage, where the rounding of the stored rows spreads near-copies apart): both
    certificates fail and the exact fp32 kernels decide.  Host buffers, device buffers and the packed (sharded) output."""
    import torch
    from nexus_b200 import FlatIndex, merge_packed
    if store == "f32":
        c, x = _near_duplicates(30000, 1000, 64, nq, 2e-5, seed=8)          # <= 16 failing queries: K2 (stream), more: KX (tile kernel)
        c[-1] = c[-2]                                         # exact duplicates too
    else:
        rng = np.random.default_rng(8)
        bg, x = gen_inputs(30000, 64, nq, seed=8)
        best = oracle.round_to_bf16(0.9 * x + 0.01 * rng.standard_normal(x.shape).astype(np.float32))
        c = np.concatenate([bg, np.repeat(best, 400, axis=0)], axis=0)
        c = np.ascontiguousarray(c[rng.permutation(c.shape[0])])
    idx = FlatIndex(64, "ip", store)
    idx.add(c)
    if mode == "numpy":
        D, I = idx.search(x, 100)
    elif mode == "cuda":
        Dt, It = idx.search(torch.from_numpy(x).cuda(), 100)
        D, I = Dt.cpu().numpy(), It.cpu().numpy()
    else:
        idx.set_id_offset(1000)
        P = idx.search_packed(torch.from_numpy(x).cuda(), 100)
        Dt, It = merge_packed(P.unsqueeze(0), [1000], 100, "ip")
        D, I = Dt.cpu().numpy(), It.cpu().numpy() - 1000
    st = idx.stats()
    assert st["cert_retry_l2"] > 0 and st["cert_retry_l3"] > 0, st
    _check(idx, c, x, 100, store, D, I)


def test_certificate_l2_metric_and_small_batches():
    from nexus_b200 import FlatIndex
    c, x = gen_inputs(50000, 96, 32, seed=77, normalize=False)
    for store in ("f32", "bf16"):
        idx = FlatIndex(96, "l2", store)
        idx.add(c)
        D, I = idx.search(x, 20)
        st = idx.stats()
        assert st["last_path_name"] == "tstream" and st["cert_checked"] == 32, st
        cs = stored(c, store)
        ref_s, ref_i = oracle.blocked_l2_search(cs, x, 20)
        rep = oracle.check_topk(ref_s, ref_i, D, I, None, None, rtol=2e-4, atol=1e-3, tie_tol=1e-3, descending=False)
        assert rep.ok, (st, rep.messages[:5])


def test_cert_off_is_round_one_behaviour():
    """NXF_CERT=0: fixed over-fetch k' = k + max(16, k/4), no certificate, nothing read back."""
    import os
    from nexus_b200 import FlatIndex
    c, x = gen_inputs(40000, 128, 200, seed=3)
    idx = FlatIndex(128, "ip", "bf16")
    idx.add(c)
    os.environ["NXF_CERT"] = "0"
    try:
        D, I = idx.search(x, 100)
    finally:
        os.environ.pop("NXF_CERT")
    st = idx.stats()
    assert st["cert_checked"] == 0 and st["last_k_fetch"] == 128, st
    _check(idx, c, x, 100, "bf16", D, I)


def test_queries_of_the_storage_type_need_no_certificate():
    import torch
    from nexus_b200 import FlatIndex
    c, x = gen_inputs(30000, 128, 200, seed=4)
    idx = FlatIndex(128, "ip", "bf16")
    idx.add(c)
    D, I = idx.search(torch.from_numpy(x).bfloat16().cuda(), 50)
    st = idx.stats()
    assert st["cert_checked"] == 0 and st["last_k_fetch"] == 50, st
    cs, xs = oracle.round_to_bf16(c), oracle.round_to_bf16(x)
    ref_s, ref_i = oracle.blocked_ip_search(cs, xs, 50)
    rep = oracle.check_topk(ref_s, ref_i, D.cpu().numpy(), I.cpu().numpy(), cs, xs, rtol=2e-4, atol=2e-5)
    assert rep.ok, rep.messages[:5]


def test_clustered_corpus_through_k1_and_k2t():
    """one CLUSTERED corpus (oracle.make_clustered: unit centroids + noise, SURVEY 8(d) variant (ii)) through the large-batch
    and the small-batch kernels, both storage types; the retry fraction is whatever the data asks for, the results are the oracle's."""
    from nexus_b200 import FlatIndex
    c = oracle.make_clustered(300000, 128, seed=1, n_centroids=256)
    x = oracle.make_clustered(300, 128, seed=2, n_centroids=256)
    for store in ("bf16", "f32"):
        idx = FlatIndex(128, "ip", store)
        idx.add(c)
        for q in (x, x[:48]):
            D, I = idx.search(q, 100)
            _check(idx, c, q, 100, store, D, I)
        st = idx.stats()
        assert st["cert_checked"] == 348, st
