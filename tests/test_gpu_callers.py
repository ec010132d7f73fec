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
