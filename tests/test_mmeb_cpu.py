"""CPU tests of the MMEB ranking block: the oracle's numpy restatement of
eval_embedding.py:345-404 against the golden vectors the reference's own code produced (tests/golden/make_golden_mmeb.py),
and the host-side list handling of nexus_b200.mmeb (no compute calls)."""
import json

import numpy as np
import pytest

import oracle
from _common import MMEB_CASES, load_mmeb_case, mmeb_golden, ranking_agrees


@pytest.mark.parametrize("name", sorted(MMEB_CASES))
def test_oracle_global_ranking_matches_reference(name):
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case(name)
    keys, scores, ranked = oracle.mmeb_global_ranking(qry, cdict)
    assert keys == list(cdict.keys())
    np.testing.assert_allclose(scores, g[f"{name}/global/scores"], rtol=1e-5, atol=1e-6)
    ref_ranked = g[f"{name}/global/ranked"]
    assert ranked.shape == ref_ranked.shape
    for qi in range(len(qry)):
        assert ranking_agrees(g[f"{name}/global/scores"][qi], ref_ranked[qi], ranked[qi]), (name, qi)
        assert sorted(ranked[qi].tolist()) == list(range(len(keys)))          # a permutation


@pytest.mark.parametrize("name", sorted(MMEB_CASES))
def test_oracle_local_ranking_matches_reference(name):
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case(name)
    off = g[f"{name}/local/offsets"]
    res = oracle.mmeb_local_ranking(qry, cdict, gts)
    assert len(res) == len(qry)
    for qi, (sc, rk) in enumerate(res):
        ref_sc = g[f"{name}/local/scores"][off[qi]:off[qi + 1]]
        ref_rk = g[f"{name}/local/ranked"][off[qi]:off[qi + 1]]
        np.testing.assert_allclose(sc, ref_sc, rtol=1e-5, atol=1e-6)
        assert ranking_agrees(ref_sc, ref_rk, rk), (name, qi)


def test_oracle_stable_order_on_exact_ties():
    """Duplicated candidates score the same up to BLAS blocking; where the floats are equal the restatement orders them by
    ascending position (a stable sort)."""
    qry, cdict, gts = load_mmeb_case("dups")
    some = 0
    for (sc, rk), gt in zip(oracle.mmeb_local_ranking(qry, cdict, gts), gts):
        names = gt["cand_names"]
        for a, b in zip(rk[:-1], rk[1:]):
            if names[a] == names[b]:
                assert abs(sc[a] - sc[b]) <= 1e-6
                assert sc[a] != sc[b] or a < b
                some += 1
    assert some > 0


@pytest.mark.parametrize("eval_type", ["global", "local"])
def test_pred_dicts_shape_and_golden_metrics_are_top10_only(eval_type):
    """The fixture stores RankingMetrics.evaluate of the reference's complete ranking; the generator asserted that the same
    call on prediction[:10] gives the same dict, which is what allows the engine to return 10 names per query."""
    g = mmeb_golden()
    qry, cdict, gts = load_mmeb_case("small")
    preds = oracle.mmeb_pred_dicts(qry, cdict, gts, eval_type)
    assert len(preds) == len(gts)
    for p, gt in zip(preds, gts):
        assert isinstance(p["label"], list) and p["rel_scores"] is None
        n = len(cdict) if eval_type == "global" else len(gt["cand_names"])
        assert len(p["prediction"]) == n
    m = json.loads(bytes(g[f"small/{eval_type}/metrics"]).decode())
    assert set(k.split("@")[1] for k in m) == {"1", "5", "10"}
    assert 0.0 < m["hit@1"] <= m["hit@5"] <= m["hit@10"] <= 1.0
    # hit@1 recomputed from the oracle's pred_dicts (metrics.py:43-53: any label in prediction[:1])
    hit1 = np.mean([float(p["prediction"][0] in p["label"]) for p in preds])
    assert abs(hit1 - m["hit@1"]) < 1e-12


def test_lists_to_csr_and_unknown_name():
    from nexus_b200.mmeb import candidate_matrix, lists_to_csr
    qry, cdict, gts = load_mmeb_case("small")
    keys, mat = candidate_matrix(cdict)
    assert keys == list(cdict.keys()) and mat.shape == (len(keys), qry.shape[1])
    row_of = {k: i for i, k in enumerate(keys)}
    lists = [gt["cand_names"] for gt in gts]
    off, rows = lists_to_csr(lists, row_of)
    assert off[0] == 0 and off[-1] == len(rows) == sum(len(l) for l in lists)
    for i, l in enumerate(lists):
        assert [keys[r] for r in rows[off[i]:off[i + 1]]] == l
    with pytest.raises(KeyError):                    # eval_embedding.py:390 raises KeyError on an unknown candidate name
        lists_to_csr([["no_such_candidate"]], row_of)
    off, rows = lists_to_csr([], row_of)
    assert off.tolist() == [0] and rows.size == 0
