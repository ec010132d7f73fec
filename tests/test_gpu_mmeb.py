"""GPU parity of the MMEB ranking block: nexus_b200.mmeb and FlatIndex.search_subsets (kernel KS,
through the C ABI nxf_search_subsets) against the oracle's restatement of eval_embedding.py:345-404 and against the golden
vectors the reference's own code produced.  Scores within rtol 1e-5; rankings identical except among scores tied within
tolerance; exactly tied scores (duplicated candidates) in ascending position order."""
import json

import numpy as np
import pytest

import oracle
from _common import MMEB_CASES, load_mmeb_case, mmeb_golden, ranking_agrees
from _gpu_common import stored

pytestmark = pytest.mark.gpu


def subset_oracle(cs, xs, offsets, rows, k, metric="ip"):
    """numpy: per query, scores of its listed rows (fp32 on the stored values), stable best-first order, -1 padding."""
    nq = len(xs)
    D = np.full((nq, k), -np.finfo(np.float32).max if metric == "ip" else np.finfo(np.float32).max, dtype=np.float32)
    P = np.full((nq, k), -1, dtype=np.int64)
    full = []
    for i in range(nq):
        r = rows[offsets[i]:offsets[i + 1]]
        if metric == "ip":
            sc = oracle.mmeb_similarity(xs[i:i + 1], cs[r])[0] if len(r) else np.zeros(0, np.float32)
            order = np.argsort(-sc, kind="stable")
        else:
            diff = cs[r].astype(np.float64) - xs[i].astype(np.float64)
            sc = (diff * diff).sum(-1).astype(np.float32) if len(r) else np.zeros(0, np.float32)
            order = np.argsort(sc, kind="stable")
        m = min(k, len(r))
        D[i, :m] = sc[order[:m]]
        P[i, :m] = order[:m]
        full.append(sc)
    return D, P, full


def check_subsets(D, P, ref_D, ref_P, full, offsets, rtol=1e-5, atol=1e-6):
    assert D.shape == ref_D.shape and P.shape == ref_P.shape and D.dtype == np.float32 and P.dtype == np.int64
    for i in range(len(full)):
        m = min(D.shape[1], int(offsets[i + 1] - offsets[i]))
        assert np.all(P[i, m:] == -1) and np.all(D[i, m:] == ref_D[i, m:]), i          # padding
        got = P[i, :m]
        assert len(set(got.tolist())) == m and (m == 0 or (got.min() >= 0 and got.max() < len(full[i]))), i
        np.testing.assert_allclose(D[i, :m], ref_D[i, :m], rtol=rtol, atol=atol)
        np.testing.assert_allclose(D[i, :m], full[i][got], rtol=rtol, atol=atol)        # the score belongs to the position
        assert ranking_agrees(full[i], ref_P[i, :m], got, rtol=4 * rtol, atol=4 * atol), i


def random_lists(rng, nq, n, lengths, dups=False):
    offsets = np.zeros(nq + 1, dtype=np.int64)
    parts = []
    for i in range(nq):
        ln = int(lengths[i % len(lengths)])
        parts.append(rng.choice(n, size=ln, replace=dups or ln > n))
        offsets[i + 1] = offsets[i] + ln
    rows = np.concatenate(parts).astype(np.int64) if parts else np.zeros(0, np.int64)
    return offsets, rows


@pytest.mark.parametrize("store", ["f32", "bf16", "f16"])
@pytest.mark.parametrize("d,lengths,k", [
    (64, [1, 7, 0, 33, 200], 10),            # 8 lanes per row; an empty list
    (128, [50, 3, 2048, 2049], 100),         # 16 lanes per row; around the shared-memory limit
    (384, [5000, 17, 1, 900], 10),           # a list parked in global scratch
    (100, [64, 10, 300], 2048),              # odd dim (zero padded), k beyond every list
    (1536, [40, 1000, 2], 1),                # long rows, k = 1
])
def test_search_subsets_matches_oracle(store, d, lengths, k):
    from nexus_b200 import FlatIndex
    rng = np.random.default_rng(d * 7 + k)
    n, nq = 6000, 23
    c = rng.standard_normal((n, d), dtype=np.float32)
    x = rng.standard_normal((nq, d), dtype=np.float32)
    offsets, rows = random_lists(rng, nq, n, lengths)
    idx = FlatIndex(d, "ip", store)
    idx.add(c)
    D, P = idx.search_subsets(x, offsets, rows, k)
    assert idx.stats()["last_path_name"] == "subset"
    ref_D, ref_P, full = subset_oracle(stored(c, store), x, offsets, rows, k)
    # unnormalised N(0,1) vectors: |q||c| ~ d, so the fp32 summation error of a score near 0 is absolute, ~1e-7 * d
    check_subsets(D, P, ref_D, ref_P, full, offsets, atol=1e-7 * d)


def test_search_subsets_exact_ties_are_in_position_order():
    """A row listed several times scores bit-identically in every slot: best-first, ties by ascending position."""
    from nexus_b200 import FlatIndex
    rng = np.random.default_rng(3)
    c = rng.standard_normal((50, 64), dtype=np.float32)
    x = rng.standard_normal((9, 64), dtype=np.float32)
    offsets, rows = random_lists(rng, 9, 50, [30, 200, 3000], dups=True)
    idx = FlatIndex(64, "ip", "f32")
    idx.add(c)
    for k in (10, 200, 2048):
        D, P = idx.search_subsets(x, offsets, rows, k)
        ref_D, ref_P, full = subset_oracle(c, x, offsets, rows, k)
        check_subsets(D, P, ref_D, ref_P, full, offsets, atol=1e-5)
        for i in range(9):
            m = min(k, int(offsets[i + 1] - offsets[i]))
            r = rows[offsets[i]:offsets[i + 1]]
            for a, b, sa, sb in zip(P[i, :m - 1], P[i, 1:m], D[i, :m - 1], D[i, 1:m]):
                assert sa >= sb
                if r[a] == r[b]:
                    assert sa == sb and a < b, (i, a, b)


def test_search_subsets_l2_metric():
    from nexus_b200 import FlatIndex
    rng = np.random.default_rng(4)
    c = rng.standard_normal((3000, 96), dtype=np.float32)
    x = rng.standard_normal((11, 96), dtype=np.float32)
    offsets, rows = random_lists(rng, 11, 3000, [25, 400, 2500])
    idx = FlatIndex(96, "l2", "f32")
    idx.add(c)
    D, P = idx.search_subsets(x, offsets, rows, 20)
    ref_D, ref_P, full = subset_oracle(c, x, offsets, rows, 20, metric="l2")
    assert np.all(np.diff(D, axis=1) >= 0)
    for i in range(11):
        np.testing.assert_allclose(D[i], ref_D[i], rtol=2e-5, atol=1e-4)
        np.testing.assert_allclose(full[i][P[i]], ref_D[i], rtol=2e-5, atol=1e-4)


def test_search_subsets_device_resident_inputs_and_outputs():
    import torch
    from nexus_b200 import FlatIndex
    rng = np.random.default_rng(5)
    c = rng.standard_normal((4000, 256), dtype=np.float32)
    x = rng.standard_normal((64, 256), dtype=np.float32)
    offsets, rows = random_lists(rng, 64, 4000, [10, 100, 1000])
    idx = FlatIndex(256, "ip", "bf16")
    idx.add(torch.from_numpy(c).cuda())
    D, P = idx.search_subsets(torch.from_numpy(x).cuda(), torch.from_numpy(offsets).cuda(), torch.from_numpy(rows).cuda(), 10)
    assert D.is_cuda and P.is_cuda and D.dtype == torch.float32 and P.dtype == torch.int64
    ref_D, ref_P, full = subset_oracle(stored(c, "bf16"), x, offsets, rows, 10)
    check_subsets(D.cpu().numpy(), P.cpu().numpy(), ref_D, ref_P, full, offsets, atol=3e-5)
    # host lists with device queries, and the other way round
    D2, P2 = idx.search_subsets(torch.from_numpy(x).cuda(), offsets, rows, 10)
    np.testing.assert_array_equal(P2.cpu().numpy(), P.cpu().numpy())
    D3, P3 = idx.search_subsets(x, torch.from_numpy(offsets).cuda(), torch.from_numpy(rows).cuda(), 10)
    np.testing.assert_array_equal(P3, P.cpu().numpy())
    np.testing.assert_array_equal(D3, D.cpu().numpy())


def test_search_subsets_errors_are_loud():
    from nexus_b200 import FlatIndex
    rng = np.random.default_rng(6)
    c = rng.standard_normal((100, 64), dtype=np.float32)
    x = rng.standard_normal((2, 64), dtype=np.float32)
    idx = FlatIndex(64, "ip", "f32")
    idx.add(c)
    with pytest.raises(ValueError, match="outside"):
        idx.search_subsets(x, np.array([0, 2, 3]), np.array([1, 100, 5]), 3)          # row 100 does not exist
    with pytest.raises(ValueError, match="outside"):
        idx.search_subsets(x, np.array([0, 2, 3]), np.array([1, -1, 5]), 3)
    with pytest.raises(ValueError):
        idx.search_subsets(x, np.array([0, 2]), np.array([1, 2]), 3)                  # nq + 1 offsets wanted
    with pytest.raises(ValueError):
        idx.search_subsets(x, np.array([0, 3, 2]), np.array([1, 2]), 3)               # not a CSR
    with pytest.raises(ValueError):
        idx.search_subsets(x, np.array([0, 1, 2]), np.array([1, 2]), 0)               # k < 1
    import torch
    with pytest.raises(ValueError, match="CSR"):                                      # device lists are checked by the kernel
        idx.search_subsets(x, torch.tensor([0, 2, 1]).cuda(), torch.tensor([1, 2]).cuda(), 3)
    D, P = idx.search_subsets(x, np.array([0, 0, 0]), np.zeros(0, np.int64), 3)       # all lists empty: padding only
    assert np.all(P == -1) and np.all(D == -np.finfo(np.float32).max)
    D, P = idx.search_subsets(x, np.array([0, 1, 2]), np.array([7, 7]), 3)            # still usable after the errors
    assert P[:, 0].tolist() == [0, 0] and np.all(P[:, 1:] == -1)


@pytest.mark.parametrize("name", sorted(MMEB_CASES))
def test_mmeb_global_matches_reference_golden(name):
    from nexus_b200 import mmeb
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case(name)
    keys, mat = mmeb.candidate_matrix(cdict)
    D, I = mmeb.rank_global(qry, mat, k=10)
    ref_scores, ref_ranked = g[f"{name}/global/scores"], g[f"{name}/global/ranked"]
    for qi in range(len(qry)):
        np.testing.assert_allclose(D[qi], ref_scores[qi][ref_ranked[qi, :10]], rtol=1e-5, atol=1e-6)
        assert ranking_agrees(ref_scores[qi], ref_ranked[qi, :10], I[qi]), (name, qi)
    if len(keys) <= 2048:                               # the complete ranking, as the reference stores it
        D, I = mmeb.rank_global(qry, mat, k=None)
        assert I.shape == ref_ranked.shape
        for qi in range(len(qry)):
            assert ranking_agrees(ref_scores[qi], ref_ranked[qi], I[qi]), (name, qi)


@pytest.mark.parametrize("name", sorted(MMEB_CASES))
@pytest.mark.parametrize("k", [10, None])
def test_mmeb_local_matches_reference_golden(name, k):
    from nexus_b200 import mmeb
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case(name)
    keys, row_of, index = mmeb.build_candidate_index(cdict)
    lists = [gt["cand_names"] for gt in gts]
    D, P = mmeb.rank_local(qry, lists, row_of, index, k=k)
    off = g[f"{name}/local/offsets"]
    for qi in range(len(qry)):
        ref_sc = g[f"{name}/local/scores"][off[qi]:off[qi + 1]]
        ref_rk = g[f"{name}/local/ranked"][off[qi]:off[qi + 1]]
        m = min(D.shape[1], len(ref_rk))
        assert np.all(P[qi, m:] == -1)
        np.testing.assert_allclose(D[qi, :m], ref_sc[ref_rk[:m]], rtol=1e-5, atol=1e-6)
        assert ranking_agrees(ref_sc, ref_rk[:m], P[qi, :m]), (name, qi)
        if k is None:
            assert m == len(ref_rk)                   # complete ranking


def _hit_at(preds, k):
    return float(np.mean([float(any(x in p["label"] for x in p["prediction"][:k])) for p in preds]))   # metrics.py:43-53


@pytest.mark.parametrize("name", sorted(MMEB_CASES))
@pytest.mark.parametrize("eval_type", ["global", "local"])
def test_build_pred_dicts_against_oracle_and_reference_metrics(name, eval_type):
    from nexus_b200 import mmeb
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case(name)
    preds = mmeb.build_pred_dicts(qry, cdict, gts, eval_type=eval_type, k=10)
    ref = oracle.mmeb_pred_dicts(qry, cdict, gts, eval_type)
    assert len(preds) == len(ref)
    same = 0
    for p, r in zip(preds, ref):
        assert p["label"] == r["label"] and p["rel_scores"] == r["rel_scores"]
        assert len(p["prediction"]) == min(10, len(r["prediction"]))
        same += p["prediction"] == r["prediction"][:10]
    assert same >= len(ref) - 1                         # names may only differ inside a tie (none expected on these inputs)
    m = json.loads(bytes(g[f"{name}/{eval_type}/metrics"]).decode())     # RankingMetrics.evaluate of the reference's ranking
    for k in (1, 5, 10):
        assert abs(_hit_at(preds, k) - m[f"hit@{k}"]) < 1e-12, (k, m)
