"""GPU parity: the CUDA paths (through the C ABI) against the oracle and the reference's golden vectors."""
import numpy as np
import pytest

import oracle
from _common import gen_inputs, golden, load_search_case, search_case_names
from _gpu_common import FLAGS, run_and_check, stored

pytestmark = pytest.mark.gpu


# ---------------------------------------------------------------- golden vectors from the reference
@pytest.mark.parametrize("name", search_case_names())
def test_golden_fp32_default_path(name):
    """fp32 storage, default dispatch == the reference's own NumpyIPIndex results (ids exact up to ties,
    scores to 1e-5), including config 1 (1k x 100k x 384, k=10)."""
    from nexus_b200 import FlatIndex
    c, x, k, ref_s, ref_i = load_search_case(name)
    idx = FlatIndex(c.shape[1], "ip", "f32")
    idx.add(c)
    D, I = idx.search(x, k)
    n = c.shape[0]
    kk = min(k, n)
    assert (I[:, kk:] == -1).all() and (D[:, kk:] == -np.finfo(np.float32).max).all()   # Faiss padding
    rep = oracle.check_topk(ref_s, ref_i, D[:, :kk], I[:, :kk], c.astype(np.float32), x.astype(np.float32), rtol=1e-5)
    assert rep.ok, (idx.stats(), rep.messages[:8], rep)


@pytest.mark.parametrize("path", ["exact", "stream"])
@pytest.mark.parametrize("name", ["tiny_ragged", "k_gt_n", "k_one", "k100", "k1000", "odd_dim", "unnormalised_d64"])
def test_golden_forced_paths_fp32(name, path):
    from nexus_b200 import FlatIndex
    c, x, k, ref_s, ref_i = load_search_case(name)
    if path == "stream":
        x = x[:9]
        ref_s, ref_i = ref_s[:9], ref_i[:9]
    idx = FlatIndex(c.shape[1], "ip", "f32")
    idx.add(c)
    D, I = idx.search(x, k, flags=FLAGS[path])
    kk = min(k, c.shape[0])
    rep = oracle.check_topk(ref_s, ref_i, D[:, :kk], I[:, :kk], c.astype(np.float32), x.astype(np.float32), rtol=1e-5)
    assert rep.ok, (idx.stats(), rep.messages[:8], rep)


def test_known_answers_through_cabi():
    from nexus_b200 import FlatIndex
    g = golden()
    for name, k in [("identity", 3), ("duplicates", 6), ("toy", 2)]:
        c, x = g[f"kat/{name}/corpus"], g[f"kat/{name}/queries"]
        idx = FlatIndex(c.shape[1], "ip", "f32")
        idx.add(c)
        D, I = idx.search(x, k)
        rep = oracle.check_topk(g[f"kat/{name}/scores"][:, :k], g[f"kat/{name}/ids"][:, :k], D, I, c, x)
        assert rep.ok, (name, rep.messages)
    D, I = idx.search(x, 10)           # toy: 2 documents, k=10 -> 8 padded columns
    assert (I[:, 2:] == -1).all()


# ---------------------------------------------------------------- every kernel family vs the oracle
CASES = [
    # N, D, Q, k
    (3000, 64, 37, 10),
    (5000, 128, 130, 100),
    (20000, 384, 300, 10),
    (9000, 768, 257, 100),
    (4096, 96, 20, 500),
    (2500, 100, 5, 7),
]


@pytest.mark.parametrize("store", ["f32", "bf16", "f16"])
@pytest.mark.parametrize("case", CASES)
def test_exact_kernel(case, store):
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=100 + n % 97)
    run_and_check(c, x, k, store=store, path="exact", expect_path="exact")


@pytest.mark.parametrize("store", ["f32", "bf16", "f16"])
@pytest.mark.parametrize("nq", [1, 2, 3, 4])
@pytest.mark.parametrize("case", [(3000, 64, 10), (20000, 384, 100), (9000, 768, 20), (7001, 1536, 10), (2500, 100, 7)])
def test_stream_kernel(case, nq, store):
    n, d, k = case
    c, x = gen_inputs(n, d, nq, seed=200 + d)
    # round 2: every small batch that fits goes to the tensor-core streaming kernel (K2T) -- one query and fp32 storage (through its
    # bf16 hi plane, certified) included; K2 (CUDA cores, exact fp32 FMAs) stays reachable by flag and serves the exact retry level
    run_and_check(c, x, k, store=store, path="auto", expect_path="tstream")
    run_and_check(c, x, k, store=store, path="stream", expect_path="stream")


@pytest.mark.parametrize("store", ["bf16", "f16"])
@pytest.mark.parametrize("case", [
    # N, D, Q, k
    (3000, 64, 2, 10), (3000, 64, 16, 10), (5000, 128, 17, 100), (20000, 384, 33, 10), (9000, 768, 64, 100),
    (4096, 96, 100, 500), (2500, 100, 5, 7), (130, 64, 128, 200), (40000, 64, 128, 500), (7001, 1536, 32, 10),
])
def test_tensor_stream_kernel(case, store):
    """K2T (corpus as the M operand, queries resident in shared memory): default dispatch for 2 <= nq <= 128 on 16-bit storage."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=500 + n % 79)
    run_and_check(c, x, k, store=store, path="auto", rtol=1e-5, expect_path="tstream", min_recall=0.999)
    run_and_check(c, x, k, store=store, path="tstream_raw", rtol=2e-4, atol=2e-5, expect_path="tstream", min_recall=0.999)


@pytest.mark.parametrize("case", [
    # N, D, Q, k, expected default path
    (3000, 64, 1, 10, "tstream"), (3000, 64, 2, 10, "tstream"), (3000, 64, 16, 10, "tstream"), (5000, 128, 17, 100, "tstream"), (20000, 384, 33, 10, "tstream"),
    (9000, 768, 32, 100, "tstream"), (9000, 768, 64, 100, "tstream"), (9000, 768, 128, 100, "tensor"), (4096, 96, 100, 500, "tstream"), (2500, 100, 5, 7, "tstream"),
    (40000, 64, 128, 500, "tstream"), (7001, 1536, 16, 10, "tstream"),
])
def test_tensor_stream_kernel_fp32_storage(case):
    """K2T over the bf16 hi plane of fp32 rows (2 bytes per element streamed), fp32 re-score, certificate."""
    n, d, q, k, expect = case
    c, x = gen_inputs(n, d, q, seed=700 + n % 67)
    run_and_check(c, x, k, store="f32", path="auto", rtol=1e-5, expect_path=expect)
    if expect == "tstream":
        run_and_check(c, x, k, store="f32", path="tstream_raw", rtol=2e-5, atol=4e-6, expect_path="tstream", min_recall=0.999)


@pytest.mark.parametrize("store,case", [
    ("bf16", (320000, 64, 128, 500)), ("bf16", (400000, 64, 7, 10)), ("f16", (310000, 128, 33, 100)), ("f32", (330000, 64, 40, 20)),
    ("bf16", (350000, 64, 128, 1)), ("bf16", (1300000, 64, 40, 10)), ("bf16", (80000, 64, 128, 300)),
])
def test_tensor_stream_kernel_seeded(store, case):
    """Longer corpora are searched in up to three launches (one tile per CTA, 16 more, the rest); between them the seed kernel raises
    tau_g to the k'-th best over all per-CTA lists of the query: same results."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=800 + n % 61)
    idx, D, I, rep, st = run_and_check(c, x, k, store=store, path="auto", rtol=1e-5, expect_path="tstream", min_recall=0.999)
    kf = k + max(16, (k + 3) // 4)
    if q * kf >= 768:                    # the library's policy: seed when nq * k' >= 768
        assert st["kernel_launches"] >= 6    # 2 converts, seed pass, seed threshold, main pass, merge


@pytest.mark.parametrize("store", ["bf16", "f32"])
def test_short_rows_many_queries_go_to_k1(store):
    """65-128 queries on 128-byte rows of a large corpus: K1 (four epilogue groups on split accumulators) instead of K2T, whose
    per-tile filter paces the stream there; 64 queries stay on K2T."""
    n, d, k = 2_100_000, 64, 10
    c, x = gen_inputs(n, d, 128, seed=1234)
    run_and_check(c, x, k, store=store, path="auto", rtol=1e-5, expect_path="tensor", min_recall=0.999)
    run_and_check(c, x[:64], k, store=store, path="auto", rtol=1e-5, expect_path="tstream", min_recall=0.999)


def test_tensor_stream_kernel_sampled_thresholds_fallback():
    """Adversarial corpus for K2T's sample (148 CTAs, 2344 tiles of 128 rows -> every 15th tile is a sample tile): all rows that
    matter lie in sample tiles, the sampled threshold is far too high, every query must go through the fall-back launch."""
    n, d, q, k = 300000, 64, 100, 100
    rng = np.random.default_rng(6)
    c = (0.05 * rng.standard_normal((n, d))).astype(np.float32)
    x = rng.standard_normal((q, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    n_tiles = (n + 127) // 128
    stride = n_tiles // 148
    hot = np.flatnonzero((np.arange(n) // 128) % stride == 0)
    c[hot] += x[rng.integers(0, q, size=hot.size)] * rng.uniform(0.5, 1.0, size=(hot.size, 1)).astype(np.float32)
    run_and_check(c, x, k, store="bf16", path="auto", rtol=1e-5, expect_path="tstream", min_recall=0.999)
    run_and_check(c, x, k, store="f32", path="auto", rtol=1e-5, expect_path="tstream", min_recall=0.999)


def test_tensor_stream_falls_back_when_queries_do_not_fit():
    """128 queries x 1536 dims = 393 KB of resident queries: K2T is not eligible, the search is served by K1."""
    c, x = gen_inputs(6000, 1536, 128, seed=9)
    run_and_check(c, x, 10, store="bf16", path="auto", expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("store", ["bf16", "f16"])
@pytest.mark.parametrize("case", CASES)
def test_tensor_kernel_raw_scores(case, store):
    """tcgen05 GEMM + fused top-k without the re-score: must equal fp32 math on the ROUNDED queries and rows."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=300 + n % 89)
    run_and_check(c, x, k, store=store, path="tensor_raw", rtol=2e-4, atol=2e-5, expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("store", ["bf16", "f16"])
@pytest.mark.parametrize("case", CASES)
def test_tensor_kernel_rescored(case, store):
    """Default tensor path: over-fetch 2k, re-score with the fp32 queries -> fp32-exact scores on the stored rows."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=400 + n % 83)
    run_and_check(c, x, k, store=store, path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("case", CASES + [(60000, 384, 700, 100), (30000, 1024, 300, 10)])
def test_tensor_kernel_fp32_storage(case):
    """fp32 rows on tcgen05 (certified search): the candidate pass streams the rows' bf16 hi plane, k' = 4k candidates are re-scored
    in fp32 and certified.  Default dispatch above the stream kernels' batch; results to the fp32 tolerance."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=600 + n % 71)
    run_and_check(c, x, k, store="f32", path="auto", rtol=1e-5, expect_path="tensor" if q > 128 else "tstream")
    run_and_check(c, x, k, store="f32", path="tensor", rtol=1e-5, expect_path="tensor")
    # without the re-score the scores are the candidate pass itself: bf16(q) . bf16(c)
    run_and_check(c, x, k, store="f32", path="tensor_raw", rtol=2e-4, atol=2e-5, expect_path="tensor", min_recall=0.999)


def test_tensor_fp32_storage_follows_adds_and_reset():
    """The bf16 plane is built on first use and extended / rebuilt when rows are added, the storage grows or the index is reset."""
    from nexus_b200 import FlatIndex
    c, x = gen_inputs(9000, 200, 200, seed=31)
    idx = FlatIndex(200, "ip", "f32")
    idx.add(c[:2000])
    for upto in (2000, 2500, 9000):
        if idx.ntotal < upto:
            idx.add(c[idx.ntotal:upto])
        D, I = idx.search(x, 20)
        assert idx.stats()["last_path_name"] == "tensor"
        ref_s, ref_i = oracle.blocked_ip_search(c[:upto], x, 20, pad=True)
        rep = oracle.check_topk(ref_s, ref_i, D, I, c[:upto], x, rtol=1e-5)
        assert rep.ok, (upto, rep.messages[:5])
    idx.reset()
    idx.add(c[5000:6000])
    D, I = idx.search(x, 20)
    ref_s, ref_i = oracle.blocked_ip_search(c[5000:6000], x, 20, pad=True)
    rep = oracle.check_topk(ref_s, ref_i, D, I, c[5000:6000], x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]


def test_tensor_fp32_storage_l2():
    c, x = gen_inputs(20000, 128, 200, seed=41, normalize=False)
    run_and_check(c, x, 50, store="f32", path="auto", metric="l2", rtol=2e-4, atol=1e-3, expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("store,case", [
    ("bf16", (300000, 64, 300, 100)), ("bf16", (270000, 128, 700, 10)), ("f16", (300000, 64, 260, 1000)), ("f32", (280000, 64, 150, 50)),
    ("bf16", (262144 + 77, 64, 129, 300)),
])
def test_tensor_kernel_sampled_thresholds(store, case):
    """Corpora of >= 1024 tiles: K1 runs two sample passes (1/256 and 1/16 of the tiles, spread evenly), starts the main pass from
    the sampled-rank threshold, verifies the candidate counts and falls back for queries that came back short: same results."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=900 + n % 59)
    idx, D, I, rep, st = run_and_check(c, x, k, store=store, path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)
    assert st["kernel_launches"] >= 8    # 2 converts, 3 passes, verification, fall-back launch, merge


def test_tensor_kernel_sampled_thresholds_fallback():
    """An adversarial corpus: every row that matters lies in a SAMPLE tile (physical tiles 0, 16, 32, ...), so the sampled threshold is
    far too high, the main pass finds nothing and the verification must send every query through the fall-back pass."""
    n, d, q, k = 300000, 64, 200, 100
    rng = np.random.default_rng(5)
    c = (0.05 * rng.standard_normal((n, d))).astype(np.float32)
    x = rng.standard_normal((q, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    tile = np.arange(n) // 256
    hot = np.flatnonzero(tile % 16 == 0)
    c[hot] += x[rng.integers(0, q, size=hot.size)] * rng.uniform(0.5, 1.0, size=(hot.size, 1)).astype(np.float32)
    run_and_check(c, x, k, store="bf16", path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)
    run_and_check(c, x, k, store="f32", path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("path,q", [("tensor", 300), ("auto", 40)])
def test_sampled_thresholds_l2_metric(path, q):
    """L2 on a corpus long enough for the sample passes (K1 with 300 queries, K2T with 40): the 0.5*|c|^2 adjustment follows the
    permuted tile order."""
    rng = np.random.default_rng(12)
    c = rng.standard_normal((290000, 64)).astype(np.float32)
    x = rng.standard_normal((q, 64)).astype(np.float32)
    run_and_check(c, x, 30, store="bf16", path=path, metric="l2", rtol=2e-4, atol=1e-3, min_recall=0.99,
                  expect_path="tensor" if path == "tensor" else "tstream")


def test_tensor_kernel_multi_launch_batches():
    """More queries than one launch holds (148 SMs: 18 944): full launches and a shorter last one, whose query tiles are shared by more
    clusters (longer arena slices).  Checked on queries from the first, the middle and the last launch.  (Short rows on a long
    search: the default picks K1's two epilogue sets here.)"""
    from nexus_b200 import FlatIndex
    n, d, q, k = 270000, 64, 19500, 20
    c, x = gen_inputs(n, d, q, seed=321)
    idx = FlatIndex(d, "ip", "bf16")
    idx.add(c)
    D, I = idx.search(x, k)
    assert idx.stats()["last_path_name"] == "tensor"
    sel = np.r_[0:40, 9000:9040, 18930:18960, 19460:19500]
    cs = stored(c, "bf16")
    ref_s, ref_i = oracle.blocked_ip_search(cs, x[sel], k, pad=True)
    rep = oracle.check_topk(ref_s, ref_i, D[sel], I[sel], cs, x[sel], rtol=1e-5)
    assert rep.ok, rep.messages[:5]
    assert (I >= 0).all() and (I < n).all()
    assert (np.diff(D, axis=1) <= 0).all()


@pytest.mark.parametrize("n", [262144 - 256, 262144, 262144 + 1])
def test_tensor_kernel_at_the_sampling_threshold(n):
    """n_tiles = 1023 / 1024 / 1025: just below and at the size from which the sample passes run."""
    c, x = gen_inputs(n, 64, 140, seed=55 + n % 7)
    run_and_check(c, x, 100, store="bf16", path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)


def test_tensor_default_dispatch_and_bf16_recall_vs_fp32():
    """BASELINE bar: bf16 recall@k >= 0.999 against the fp32 reference on the same (bf16-valued) corpus."""
    c, x = gen_inputs(50000, 768, 512, seed=77)
    c = oracle.round_to_bf16(c)                     # the corpus IS bf16 (config 2's wording)
    idx, D, I, rep, st = run_and_check(c, x, 100, store="bf16", path="auto", expect_path="tensor", min_recall=0.999)
    assert st["last_k_fetch"] == 200 and st["cert_checked"] == 512 and st["cert_retry_l2"] == 0      # k' = 2k, every query certified


def test_multiple_adds_and_reset():
    c, x = gen_inputs(6000, 128, 50, seed=5)
    idx, D, I, rep, st = run_and_check(c, x, 10, store="bf16", path="auto", add_chunks=3)
    idx.reset()
    assert idx.ntotal == 0
    D, I = idx.search(x, 5)
    assert (I == -1).all()
    idx.add(c[:100])
    D2, I2 = idx.search(x, 5)
    assert I2.max() < 100 and (I2 >= 0).all()


def test_reconstruct_roundtrip():
    from nexus_b200 import FlatIndex
    c, _ = gen_inputs(1000, 100, 1, seed=9)
    for store in ["f32", "bf16", "f16"]:
        idx = FlatIndex(100, "ip", store)
        idx.add(c)
        np.testing.assert_array_equal(idx.reconstruct_n(), stored(c, store))
        np.testing.assert_array_equal(idx.reconstruct_n(10, 5), stored(c, store)[10:15])


@pytest.mark.parametrize("path,store", [("exact", "f32"), ("stream", "f32"), ("tensor", "bf16"), ("stream", "bf16")])
def test_l2_metric(path, store):
    rng = np.random.default_rng(4)
    c = rng.standard_normal((5000, 64)).astype(np.float32)
    x = rng.standard_normal((4 if path == "stream" else 70, 64)).astype(np.float32)
    run_and_check(c, x, 20, store=store, path=path, metric="l2", rtol=2e-4, atol=1e-3, min_recall=0.99)


def test_torch_cuda_in_out_and_id_offset():
    import torch
    from nexus_b200 import FlatIndex
    c, x = gen_inputs(8000, 64, 300, seed=31, normalize=False)
    ct, xt = torch.from_numpy(c).cuda(), torch.from_numpy(x).cuda()
    idx = FlatIndex(64, "ip", "f32")
    idx.add(ct)                                     # device rows: modeling.py:195
    D, I = idx.search(xt, 50)                       # device queries -> device results: modeling.py:197
    assert D.is_cuda and I.is_cuda and D.dtype == torch.float32 and I.dtype == torch.int64
    ref_s, ref_i = oracle.blocked_ip_search(c, x, 50)
    rep = oracle.check_topk(ref_s, ref_i, D.cpu().numpy(), I.cpu().numpy(), c, x)
    assert rep.ok, rep.messages[:5]
    idx.set_id_offset(1000)
    D2, I2 = idx.search(xt[:8], 5)
    np.testing.assert_array_equal(I2.cpu().numpy(), ref_i[:8, :5] + 1000)
    # bf16 tensors straight from an encoder
    idx2 = FlatIndex(64, "ip", "bf16")
    idx2.add(ct.bfloat16())
    D3, I3 = idx2.search(xt.bfloat16(), 10)
    cs, xs = oracle.round_to_bf16(c), oracle.round_to_bf16(x)
    r_s, r_i = oracle.blocked_ip_search(cs, xs, 10)
    rep = oracle.check_topk(r_s, r_i, D3.cpu().numpy(), I3.cpu().numpy(), cs, xs, rtol=2e-4, atol=2e-5)
    assert rep.ok, rep.messages[:5]


def test_sharded_equals_unsharded_single_gpu():
    """KAT-5: G logical shards + merge kernel == one index (multi-GPU logic on one device)."""
    import torch
    from nexus_b200 import FlatIndex, merge_topk, shard_bounds
    c, x = gen_inputs(30000, 128, 200, seed=41)
    full = FlatIndex(128, "ip", "f32")
    full.add(c)
    D0, I0 = full.search(x, 100)
    for G in (2, 4, 8):
        Ds, Is = [], []
        for lo, hi in shard_bounds(c.shape[0], G):
            s = FlatIndex(128, "ip", "f32")
            s.set_id_offset(lo)
            s.add(c[lo:hi])
            d_, i_ = s.search(torch.from_numpy(x).cuda(), 100)
            Ds.append(d_)
            Is.append(i_)
        D, I = merge_topk(torch.stack(Ds), torch.stack(Is), 100)
        rep = oracle.check_topk(D0, I0, D.cpu().numpy(), I.cpu().numpy(), c, x)
        assert rep.ok, (G, rep.messages[:5])
        assert rep.exact_id_rows >= 0.99 * x.shape[0]


def test_retrieval_index_search_signatures():
    """The reference-shaped functions (utils.py:148,194) end to end, numpy in / numpy out."""
    import nexus_b200
    c, x, k, ref_s, ref_i = load_search_case("bge_small_slice")
    idx = nexus_b200.index(corpus_embeddings=c, device="cuda:0")
    s, i = nexus_b200.search(idx, k=k, query_embeddings=x)
    assert s.dtype == np.float32 and i.dtype == np.int64 and s.shape == (70, 10)
    rep = oracle.check_topk(ref_s, ref_i, s, i, c, x)
    assert rep.ok, rep.messages[:5]
    with pytest.raises(ValueError):
        nexus_b200.index(index_factory="IVF100,PQ8", corpus_embeddings=c)


def test_errors_are_loud():
    from nexus_b200 import FlatIndex
    idx = FlatIndex(16, "ip", "f32")
    with pytest.raises(ValueError):
        idx.add(np.zeros((3, 8), np.float32))
    idx.add(np.ones((3, 16), np.float32))
    with pytest.raises(ValueError):
        idx.search(np.zeros((1, 16), np.float32), 0)
    with pytest.raises(ValueError):
        idx.search(np.zeros((1, 16), np.float32), 5000)
    with pytest.raises(Exception):
        idx.search(np.zeros((9, 16), np.float32), 2, flags=FLAGS["tensor"] | FLAGS["exact"])   # two forced paths


@pytest.fixture
def k1_sets(request):
    """NXF_K1_SETS for the duration of one test (the library reads it per search)."""
    import os
    old = os.environ.get("NXF_K1_SETS")
    os.environ["NXF_K1_SETS"] = str(request.param)
    yield request.param
    if old is None:
        os.environ.pop("NXF_K1_SETS", None)
    else:
        os.environ["NXF_K1_SETS"] = old


@pytest.mark.parametrize("k1_sets", [1, 2], indirect=True)
@pytest.mark.parametrize("path", ["tensor", "tensor_1cta"])
@pytest.mark.parametrize("store,case", [
    ("bf16", (3000, 64, 37, 10)), ("f16", (5000, 128, 130, 100)), ("bf16", (20000, 384, 300, 10)), ("bf16", (9000, 512, 257, 100)),
    ("bf16", (4096, 96, 20, 500)), ("f32", (6000, 64, 140, 20)),
])
def test_tensor_kernel_epilogue_sets(k1_sets, path, store, case):
    """K1 with one and with two epilogue warp sets (the second set filters the other TMEM buffer into its own per-row lists; the
    default picks it for short rows on long searches only, so small tests would never run it): same results either way."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=1300 + d)
    run_and_check(c, x, k, store=store, path=path, rtol=1e-5, expect_path="tensor", min_recall=0.999)


@pytest.mark.parametrize("k1_sets", [2], indirect=True)
@pytest.mark.parametrize("store,case", [("bf16", (300000, 64, 300, 100)), ("f16", (300000, 64, 260, 1000)), ("bf16", (270000, 128, 700, 10))])
def test_tensor_kernel_two_sets_sampled_thresholds(k1_sets, store, case):
    """Two epilogue sets through the sampled-threshold schedule (estimate, two sample passes, main, verification, fall-back)."""
    n, d, q, k = case
    c, x = gen_inputs(n, d, q, seed=1400 + n % 59)
    idx, D, I, rep, st = run_and_check(c, x, k, store=store, path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)
    assert st["kernel_launches"] >= 8


@pytest.mark.parametrize("k1_sets", [2], indirect=True)
def test_tensor_kernel_two_sets_fallback(k1_sets):
    n, d, q, k = 300000, 64, 200, 100
    rng = np.random.default_rng(6)
    c = (0.05 * rng.standard_normal((n, d))).astype(np.float32)
    x = rng.standard_normal((q, d)).astype(np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    hot = np.flatnonzero((np.arange(n) // 256) % 16 == 0)      # every row that matters lies in a sample tile: the fall-back must run
    c[hot] += x[rng.integers(0, q, size=hot.size)] * rng.uniform(0.5, 1.0, size=(hot.size, 1)).astype(np.float32)
    run_and_check(c, x, k, store="bf16", path="tensor", rtol=1e-5, expect_path="tensor", min_recall=0.999)


def test_two_streams_one_handle_no_caller_sync():
    """SURVEY.md 8(b) "concurrent search on distinct streams": one handle, two non-blocking streams, interleaved device-in /
    device-out searches of different shapes (K1 and K2T carve the shared scratch arena differently) with no synchronisation by the
    caller between them.  The library orders the calls on the device (include/nexus_flat.h "Thread-safety and streams"); every
    result is compared with the oracle."""
    import torch
    from nexus_b200 import FlatIndex
    c, xa = gen_inputs(60000, 128, 300, seed=77)
    _, xb = gen_inputs(8, 128, 16, seed=78)
    idx = FlatIndex(128, "ip", "bf16")
    idx.add(c)
    cs = oracle.round_to_bf16(c)
    ref_a = oracle.blocked_ip_search(cs, xa, 50)
    ref_b = oracle.blocked_ip_search(cs, xb, 10)
    qa, qb = torch.from_numpy(xa).cuda(), torch.from_numpy(xb).cuda()
    sa, sb = torch.cuda.Stream(), torch.cuda.Stream()
    torch.cuda.synchronize()
    outs = []
    for it in range(50):
        with torch.cuda.stream(sa):
            Da, Ia = idx.search(qa, 50)
        with torch.cuda.stream(sb):
            Db, Ib = idx.search(qb, 10)
        outs.append((Da, Ia, Db, Ib))
    torch.cuda.synchronize()
    for Da, Ia, Db, Ib in outs:
        rep = oracle.check_topk(ref_a[0], ref_a[1], Da.cpu().numpy(), Ia.cpu().numpy(), cs, xa)
        assert rep.ok, rep.messages[:5]
        rep = oracle.check_topk(ref_b[0], ref_b[1], Db.cpu().numpy(), Ib.cpu().numpy(), cs, xb)
        assert rep.ok, rep.messages[:5]
    # add on one stream, search on another, no sync: the search must see the converted rows
    idx2 = FlatIndex(128, "ip", "bf16")
    ct = torch.from_numpy(c).cuda()
    torch.cuda.synchronize()
    with torch.cuda.stream(sa):
        idx2.add(ct)
    with torch.cuda.stream(sb):
        D2, I2 = idx2.search(qa, 50)
    torch.cuda.synchronize()
    rep = oracle.check_topk(ref_a[0], ref_a[1], D2.cpu().numpy(), I2.cpu().numpy(), cs, xa)
    assert rep.ok, rep.messages[:5]
