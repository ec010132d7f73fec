"""GPU parity of the callers right after the search: rank of the positive among the top-k, through the
C ABI (nxf_positive_rank), against the oracle's restatement of modeling.py:199-205.  Bit-exact (integer / boolean)."""
import numpy as np
import pytest

import oracle

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,k", [(1, 1), (5, 7), (64, 10), (333, 100), (4096, 50), (17, 1000)])
def test_positive_rank_matches_reference_sort(B, k):
    import torch
    from nexus_b200 import positive_rank
    rng = np.random.default_rng(B * 31 + k)
    topk = np.sort(rng.standard_normal((B, k)).astype(np.float32), axis=1)[:, ::-1].copy()
    pos = rng.standard_normal(B).astype(np.float32)
    for b in range(0, B, 3):                       # ties with a top-k score
        pos[b] = topk[b, rng.integers(0, k)]
    if B > 2:
        pos[1] = topk[1].max() + 1                 # rank 0
        pos[2] = topk[2].min() - 1                 # behind all k
    rank, pred = positive_rank(torch.from_numpy(pos).cuda(), torch.from_numpy(topk).cuda())
    ref_rank, ref_pred = oracle.positive_pred(pos, topk)
    np.testing.assert_array_equal(rank.cpu().numpy(), ref_rank)
    np.testing.assert_array_equal(pred.cpu().numpy(), ref_pred)
    only = positive_rank(torch.from_numpy(pos).cuda().unsqueeze(-1), torch.from_numpy(topk).cuda(), return_pred=False)
    np.testing.assert_array_equal(only.cpu().numpy(), ref_rank)


def test_positive_rank_after_a_search():
    """eval_step end to end: search the item table, score the positive, rank it (modeling.py:190-205)."""
    import torch
    from nexus_b200 import FlatIndex, positive_rank
    gen = torch.Generator().manual_seed(5)
    items = torch.randn(20000, 64, generator=gen)
    q = torch.randn(256, 64, generator=gen)
    target = torch.randint(0, items.shape[0], (256,), generator=gen)
    q = q + 2.0 * items[target]                    # the positive scores high for most rows
    index = FlatIndex(64, "ip", store_dtype="f32", device=0)
    index.add(items.numpy())
    scores, ids = index.search(q.numpy(), 20)
    pos = (q * items[target]).sum(-1)
    rank, pred = positive_rank(pos.cuda(), torch.from_numpy(scores).cuda())
    ref_rank, ref_pred = oracle.positive_pred(pos.numpy(), scores)
    np.testing.assert_array_equal(rank.cpu().numpy(), ref_rank)
    np.testing.assert_array_equal(pred.cpu().numpy(), ref_pred)
    assert (ref_rank < 20).mean() > 0.5


def test_positive_rank_errors_are_loud():
    import torch
    from nexus_b200 import positive_rank
    with pytest.raises(ValueError):
        positive_rank(torch.zeros(3), torch.zeros(3, 4))
    with pytest.raises(ValueError):
        positive_rank(torch.zeros(2).cuda(), torch.zeros(3, 4).cuda())


def test_recsys_item_index_flow_from_item_vectors_pt(tmp_path):
    """Nexus/inference/utils.py:8-25 (gen_item_index) and recommendation/base.py:114-117,182 on the shim: item_vectors.pt ->
    IndexIVFFlat(quantizer, d, 100, IP) -> train/add -> write_index -> read_index -> nprobe -> search -> item_ids[I].
    The exact scan must return the true top-k (what IVF returns when every list is probed)."""
    import torch
    from nexus_b200 import faiss_shim as F
    rng = np.random.default_rng(21)
    items = rng.standard_normal((5000, 64)).astype(np.float32)
    item_ids = rng.permutation(100000)[:5000].astype(np.int64)
    torch.save({"item_vectors": torch.from_numpy(items), "item_ids": torch.from_numpy(item_ids)}, tmp_path / "item_vectors.pt")
    rows, ids = F.load_embedding_cache(str(tmp_path / "item_vectors.pt"))
    quantizer = F.IndexFlatIP(rows.shape[1])
    index = F.IndexIVFFlat(quantizer, rows.shape[1], 100, F.METRIC_INNER_PRODUCT)
    index.train(rows)
    index.add(rows)
    F.write_index(index, str(tmp_path / "item.index"))
    served = F.read_index(str(tmp_path / "item.index"))
    served.nprobe = 10
    users = rng.standard_normal((128, 64)).astype(np.float32)
    D, I = served.search(users, 50)
    ref_s, ref_i = oracle.blocked_ip_search(items, users, 50)
    rep = oracle.check_topk(ref_s, ref_i, D, I, items, users, rtol=1e-5, atol=1e-5)
    assert rep.ok, rep.messages[:5]
    np.testing.assert_array_equal(ids[I][:, 0], item_ids[ref_i[:, 0]])
    idx2, ids2 = F.index_from_cache(str(tmp_path / "item_vectors.pt"))
    D2, I2 = idx2.search(users, 50)
    np.testing.assert_array_equal(I2, I)


def test_read_index_of_a_faiss_native_flat_file(tmp_path):
    """serve/inference_service_text_retrieval.py:35,70: faiss.read_index(path) at import, index.search(q[None, :], k=20) per request --
    on a file in Faiss's native flat layout (written here by write_index_faiss_flat; parity of the layout itself is unpinned)."""
    from nexus_b200 import faiss_shim as F
    rng = np.random.default_rng(33)
    docs = rng.standard_normal((3000, 384)).astype(np.float32)
    docs /= np.linalg.norm(docs, axis=1, keepdims=True)
    index = F.IndexFlatIP(384)
    index.add(docs)
    F.write_index_faiss_flat(index, str(tmp_path / "doc.index"))
    served = F.read_index(str(tmp_path / "doc.index"))
    assert served.ntotal == 3000 and served.d == 384
    q = rng.standard_normal(384).astype(np.float32)
    D, I = served.search(q[None, :], 20)
    ref_s, ref_i = oracle.blocked_ip_search(docs, q[None, :], 20)
    rep = oracle.check_topk(ref_s, ref_i, D, I, docs, q[None, :], rtol=1e-5)
    assert rep.ok, rep.messages[:5]


# ------------------------------------------------------------------------------------------------------------------
# GPU half of "run the reference's callers on the engine" (rows a7 / a8 / a9).  /root/reference does not exist on the GPU box, so
# the callers' statements are replayed here one for one, with `faiss` = the shim, on the real library; the CPU half
# (tests/test_reference_callers_cpu.py) runs the reference's own function objects on an oracle-backed fake of the library.
# ------------------------------------------------------------------------------------------------------------------
def _replay_eval_step(faiss, scorer, query_vec, pos_vec, k, item_vectors, device_index):
    """Nexus/training/embedder/recommendation/modeling.py:179-205, statement for statement."""
    import torch
    if scorer == "InnerProductScorer":
        pos_scores = (query_vec * pos_vec).sum(-1)                                  # :181 (modules/score.py InnerProductScorer)
    elif scorer == "EuclideanScorer":
        pos_scores = -((query_vec - pos_vec) ** 2).sum(-1)                          # :181 (modules/score.py:53-59: minus the squared distance)
    else:
        pos_scores = torch.nn.functional.cosine_similarity(query_vec, pos_vec, dim=-1)
    res = faiss.StandardGpuResources()                                              # :182
    if scorer == "InnerProductScorer":
        index_flat = faiss.IndexFlatIP(item_vectors.shape[-1])                      # :184
    elif scorer == "EuclideanScorer":
        index_flat = faiss.IndexFlatL2(item_vectors.shape[-1])                      # :186
    else:
        item_vectors = torch.nn.functional.normalize(item_vectors, p=2, dim=-1)     # :188
        pos_vec = torch.nn.functional.normalize(pos_vec, p=2, dim=-1)               # :189
        index_flat = faiss.IndexFlatIP(item_vectors.shape[-1])                      # :190
    gpu_index = device_index                                                        # :193
    index_flat = faiss.index_cpu_to_gpu(res, gpu_index, index=index_flat)           # :194 (index= by keyword)
    index_flat.add(item_vectors)                                                    # :195
    topk_scores, topk_indices = index_flat.search(query_vec, k)                     # :197
    if pos_scores.dim() < topk_scores.dim():
        pos_scores = pos_scores.unsqueeze(-1)                                       # :200-201
    all_scores = torch.cat([pos_scores, topk_scores], dim=-1)                       # :202
    _, indice = torch.sort(all_scores, dim=-1, descending=True, stable=True)        # :204
    pred = indice[:, :k] == 0                                                       # :205
    return pred, topk_scores, topk_indices, pos_scores, index_flat


@pytest.mark.parametrize("scorer", ["InnerProductScorer", "EuclideanScorer", "CosineScorer"])
def test_eval_step_replayed_on_the_shim(scorer):
    """Row a9 + N3: 20 consecutive eval batches against the same item table.  Results: top-k vs the oracle, ``pred`` vs the oracle's
    restatement AND vs nxf_positive_rank; the item table is converted and uploaded ONCE (identity hit for IP / L2, checksum hit for
    the Cosine scorer's freshly normalised tensor)."""
    import torch
    from nexus_b200 import faiss_shim as F, positive_rank
    F._item_cache.clear()
    for kname in F.cache_stats:
        F.cache_stats[kname] = 0
    gen = torch.Generator().manual_seed(11)
    N, D, B, k = 50000, 64, 128, 20
    item_vectors = torch.randn(N, D, generator=gen).cuda()
    items_np = item_vectors.cpu().numpy()
    if scorer == "CosineScorer":
        items_np = torch.nn.functional.normalize(item_vectors, p=2, dim=-1).cpu().numpy()
    impls, rows_added = set(), None
    for batch in range(20):
        target = torch.randint(0, N, (B,), generator=gen)
        query_vec = (torch.randn(B, D, generator=gen) + 1.2 * item_vectors[target].cpu()).cuda()
        pos_vec = (item_vectors[target].cpu() + 0.05 * torch.randn(B, D, generator=gen)).cuda()
        pred, top_s, top_i, pos_s, index_flat = _replay_eval_step(F, scorer, query_vec, pos_vec, k, item_vectors, 0)
        assert top_s.is_cuda and top_i.is_cuda and top_i.dtype == torch.int64 and pred.dtype == torch.bool
        impls.add(id(index_flat._impl))
        rows_added = index_flat._impl.stats()["rows_added"]
        q_np = query_vec.cpu().numpy()
        if scorer == "EuclideanScorer":
            ref_s, ref_i = oracle.blocked_l2_search(items_np, q_np, k)
            rep = oracle.check_topk(ref_s, ref_i, top_s.cpu().numpy(), top_i.cpu().numpy(), None, None, rtol=1e-5, atol=1e-4, tie_tol=1e-3, descending=False)
        else:
            ref_s, ref_i = oracle.blocked_ip_search(items_np, q_np, k)
            rep = oracle.check_topk(ref_s, ref_i, top_s.cpu().numpy(), top_i.cpu().numpy(), items_np, q_np, rtol=1e-5, atol=1e-5)
        assert rep.ok, (batch, rep.messages[:5])
        rank_ref, pred_ref = oracle.positive_pred(pos_s.reshape(-1).cpu().numpy(), top_s.cpu().numpy())
        np.testing.assert_array_equal(pred.cpu().numpy(), pred_ref)
        rank, pred_k = positive_rank(pos_s, top_s)
        np.testing.assert_array_equal(pred_k.cpu().numpy(), pred_ref)
    assert len(impls) == 1 and rows_added == N, (len(impls), rows_added)              # N3: the shard was built once
    want = "checksum" if scorer == "CosineScorer" else "identity"
    assert F.cache_stats["miss"] == 1 and F.cache_stats[want] == 19, F.cache_stats
    # a changed table (in-place update bumps _version; new contents change the checksum) must NOT hit
    item_vectors.mul_(1.5)
    _, _, _, _, index2 = _replay_eval_step(F, scorer, query_vec, pos_vec, k, item_vectors, 0)
    if scorer != "CosineScorer":                                                       # normalisation undoes a uniform scale: same rows
        assert id(index2._impl) not in impls
    # a further add on an object holding a shared shard takes a private copy first
    index2.add(item_vectors[:10])
    assert index2.ntotal == N + 10 and not index2._shared
    F._item_cache.clear()


def test_index_search_replayed_with_all_gpus_branch():
    """Row a7: Nexus/evaluation/text_retrieval/utils.py:176-189 + :221-229 with `faiss` = the shim, statement for statement:
    index_factory -> GpuMultipleClonerOptions(shard, useFloat16) -> index_cpu_to_all_gpus(index, co) -> train -> add -> 32-row
    query chunks.  With several visible devices the object is a ShardedFlatIndex (what utils.py:180-183 asks Faiss for)."""
    import torch
    from nexus_b200 import faiss_shim as faiss
    from nexus_b200.sharded import ShardedFlatIndex
    c, x = gen_inputs_local(40000, 96, 100, seed=5)
    corpus_embeddings = c.astype(np.float32)                                         # :169
    faiss_index = faiss.index_factory(corpus_embeddings.shape[-1], "Flat", faiss.METRIC_INNER_PRODUCT)   # :176
    co = faiss.GpuMultipleClonerOptions()                                            # :180
    co.shard = True                                                                  # :181
    co.useFloat16 = True                                                             # :182
    faiss_index = faiss.index_cpu_to_all_gpus(faiss_index, co)                       # :183
    faiss_index.train(corpus_embeddings)                                             # :188
    faiss_index.add(corpus_embeddings)                                               # :189
    assert faiss_index.ntotal == 40000
    if torch.cuda.device_count() > 1:
        assert isinstance(faiss_index._impl, ShardedFlatIndex) and len(faiss_index._impl.shards) == torch.cuda.device_count()
    all_scores, all_indices = [], []
    for i in range(0, len(x), 32):                                                   # :221-223
        scores, indices = faiss_index.search(x[i:i + 32].astype(np.float32), 50)     # :224
        all_scores.append(scores)
        all_indices.append(indices)
    all_scores = np.concatenate(all_scores, axis=0)                                  # :228-229
    all_indices = np.concatenate(all_indices, axis=0)
    cs = oracle.round_to_f16(c)                                                      # useFloat16: the stored values
    ref_s, ref_i = oracle.blocked_ip_search(cs, x, 50)
    rep = oracle.check_topk(ref_s, ref_i, all_scores, all_indices, cs, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]
    # the same object through the drop-in search(): ONE call for the whole matrix (_LazyFlat is dispatched, not looped)
    import nexus_b200
    s2, i2 = nexus_b200.search(faiss_index, k=50, query_embeddings=x)
    rep = oracle.check_topk(ref_s, ref_i, s2, i2, cs, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]


def gen_inputs_local(n, d, q, seed):
    from _common import gen_inputs
    return gen_inputs(n, d, q, seed)


@pytest.mark.parametrize("ignore", [False, True])
def test_dense_retriever_body_replayed(ignore):
    """Row a8: searcher.py:149-159 -- index(), search(), then the nested loop -- against unpack_results on the same arrays, with
    k larger than the corpus so that -1 padding is in play (searcher.py:156)."""
    import nexus_b200
    c, x = gen_inputs_local(300, 48, 41, seed=9)
    corpus_ids = [f"doc-{i}" for i in range(300)]
    queries_ids = [f"q-{i}" if i % 4 else f"doc-{i}" for i in range(41)]
    x[0] = c[0]                                                                      # query "doc-0" retrieves document "doc-0"
    faiss_index = nexus_b200.index(corpus_embeddings=c, device="cuda:0")             # :149
    all_scores, all_indices = nexus_b200.search(query_embeddings=x, faiss_index=faiss_index, k=1000)   # :150 (search_top_k default)
    assert all_scores.shape == (41, 1000) and (all_indices[:, 300:] == -1).all()
    results = {}
    for idx, (scores, indices) in enumerate(zip(all_scores, all_indices)):           # :152-159
        results[queries_ids[idx]] = {}
        for score, indice in zip(scores, indices):
            if indice != -1:
                if ignore and corpus_ids[indice] == queries_ids[idx]:
                    continue
                results[queries_ids[idx]][corpus_ids[indice]] = float(score)
    got = nexus_b200.unpack_results(queries_ids, corpus_ids, all_scores, all_indices, ignore)
    assert got == results and all(len(v) == (299 if (ignore and q.startswith("doc-")) else 300) for q, v in got.items())
    ref_s, ref_i = oracle.blocked_ip_search(c, x, 300)
    rep = oracle.check_topk(ref_s, ref_i, all_scores[:, :300], all_indices[:, :300], c, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]
    if ignore:
        assert "doc-0" not in got["doc-0"]


def test_index_cpu_to_gpu_reuses_a_resident_shard():
    """recommendation/base.py:114-116 does read_index + index_cpu_to_gpu -- the populated index must not be uploaded twice."""
    from nexus_b200 import faiss_shim as F
    rng = np.random.default_rng(2)
    rows = rng.standard_normal((2000, 32)).astype(np.float32)
    idx = F.IndexFlatIP(32)
    idx.add(rows)
    impl = idx._impl
    same = F.index_cpu_to_gpu(F.StandardGpuResources(), 0, idx)
    assert same is idx and same._impl is impl and impl.stats()["rows_added"] == 2000
    co = F.GpuClonerOptions()
    co.useFloat16 = True
    half = F.index_cpu_to_gpu(F.StandardGpuResources(), 0, idx, co)
    assert half is not idx and half._impl.store_dtype == "f16" and half.ntotal == 2000
    q = rng.standard_normal((5, 32)).astype(np.float32)
    D, I = half.search(q, 10)
    cs = oracle.round_to_f16(rows)
    ref_s, ref_i = oracle.blocked_ip_search(cs, q, 10)
    rep = oracle.check_topk(ref_s, ref_i, D, I, cs, q, rtol=1e-5)
    assert rep.ok, rep.messages[:5]
