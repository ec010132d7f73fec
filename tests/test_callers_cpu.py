"""Host logic right after the search: the array-native result unpack against the reference's nested loop, and
the oracle of the positive-rank epilogue (N3) against the reference's own torch lines."""
import numpy as np
import pytest

import oracle
from nexus_b200.retrieval import unpack_results


def _case(q, k, n, seed, pad_rows=()):
    rng = np.random.default_rng(seed)
    scores = np.sort(rng.standard_normal((q, k)).astype(np.float32), axis=1)[:, ::-1].copy()
    ids = np.stack([rng.choice(n, size=k, replace=False) for _ in range(q)]).astype(np.int64)
    for r in pad_rows:
        cut = rng.integers(0, k)
        ids[r, cut:] = -1
        scores[r, cut:] = -np.finfo(np.float32).max
    return scores, ids


@pytest.mark.parametrize("ignore", [False, True])
def test_unpack_results_equals_the_reference_loop(ignore):
    q, k, n = 37, 20, 500
    scores, ids = _case(q, k, n, seed=1, pad_rows=(0, 5, 36))
    corpus_ids = [f"doc{i}" for i in range(n)]
    queries_ids = [f"doc{i}" if i % 3 == 0 else f"q{i}" for i in range(q)]   # some query ids ARE corpus ids
    ids[3, 2] = 3                                                             # ... and appear in their own results
    ids[6, 0] = 6
    got = unpack_results(queries_ids, corpus_ids, scores, ids, ignore_identical_ids=ignore)
    ref = oracle.unpack_results_loop(queries_ids, corpus_ids, scores, ids, ignore_identical_ids=ignore)
    assert got == ref
    assert all(type(v) is float for d in got.values() for v in d.values())
    assert [list(got[qid]) for qid in queries_ids] == [list(ref[qid]) for qid in queries_ids]   # insertion order too
    if ignore:
        assert "doc3" not in got["doc3"] and "doc6" not in got["doc6"]


def test_unpack_results_all_padding_and_errors():
    scores = np.full((2, 4), -np.finfo(np.float32).max, np.float32)
    ids = np.full((2, 4), -1, np.int64)
    assert unpack_results(["a", "b"], ["x"], scores, ids) == {"a": {}, "b": {}}
    with pytest.raises(ValueError):
        unpack_results(["a"], ["x"], scores, ids)
    with pytest.raises(ValueError):
        unpack_results(["a", "b"], ["x"], scores, ids[:, :3])


def test_positive_pred_oracle_equals_the_reference_torch_lines():
    import torch
    rng = np.random.default_rng(3)
    B, k = 64, 17
    topk = np.sort(rng.standard_normal((B, k)).astype(np.float32), axis=1)[:, ::-1].copy()
    pos = rng.standard_normal(B).astype(np.float32)
    pos[:8] = topk[:8, 4]            # ties with a top-k score: the positive sorts first (stable, it is element 0)
    pos[8] = topk[8].max() + 1       # rank 0
    pos[9] = topk[9].min() - 1       # behind all k: no True in the row
    rank, pred = oracle.positive_pred(pos, topk)
    # the reference's lines, verbatim semantics (modeling.py:199-204)
    pos_t, topk_t = torch.from_numpy(pos), torch.from_numpy(topk)
    pos_t = pos_t.unsqueeze(-1)
    all_scores = torch.cat([pos_t, topk_t], dim=-1)
    _, indice = torch.sort(all_scores, dim=-1, descending=True, stable=True)
    ref_pred = (indice[:, :k] == 0).numpy()
    np.testing.assert_array_equal(pred, ref_pred)
    np.testing.assert_array_equal(rank, (topk > pos[:, None]).sum(axis=1))
    assert rank[8] == 0 and rank[9] == k and not pred[9].any()
    assert (rank[:8] == 4).all()
