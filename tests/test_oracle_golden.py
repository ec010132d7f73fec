"""CPU: pin the oracle against golden vectors produced by the REFERENCE's own code
(tests/golden/make_golden.py ran Nexus/evaluation/text_retrieval/utils.py and the embedders'
pooling functions unmodified)."""
import numpy as np
import pytest

import oracle
from _common import golden, load_search_case, search_case_names

SMALL = [n for n in search_case_names() if n != "c1_bge_small"]


@pytest.mark.parametrize("name", SMALL)
def test_numpy_ip_search_matches_reference_bitwise(name):
    c, x, k, ref_s, ref_i = load_search_case(name)
    s, i = oracle.reference_chunked_search(lambda q, kk: oracle.numpy_ip_search(c, q, kk), x, k)
    assert s.shape == ref_s.shape and i.shape == ref_i.shape          # truncation when k > N (utils.py:23)
    assert s.dtype == np.float32 and i.dtype == np.int64
    np.testing.assert_array_equal(i, ref_i)
    np.testing.assert_array_equal(s, ref_s)


@pytest.mark.parametrize("name", search_case_names())
def test_blocked_restatement_matches_reference(name):
    c, x, k, ref_s, ref_i = load_search_case(name)
    s, i = oracle.blocked_ip_search(c, x, k, block_rows=4096 if c.shape[0] > 5000 else 97, pad=False)
    assert s.shape == ref_s.shape
    rep = oracle.check_topk(ref_s, ref_i, s, i, c.astype(np.float32), x.astype(np.float32), rtol=1e-5)
    assert rep.ok, rep.messages
    assert rep.recall == 1.0 or rep.tie_swaps > 0


def test_faiss_padding_convention():
    c, x, k, ref_s, ref_i = load_search_case("k_gt_n")
    s, i = oracle.blocked_ip_search(c, x, k, pad=True)
    assert s.shape == (x.shape[0], k)
    n = c.shape[0]
    np.testing.assert_array_equal(i[:, :n], ref_i)
    assert (i[:, n:] == -1).all() and (s[:, n:] == -np.finfo(np.float32).max).all()


def test_known_answers():
    g = golden()
    s, i = oracle.numpy_ip_search(g["kat/identity/corpus"], g["kat/identity/queries"], 3)
    assert i[:, 0].tolist() == [3, 7, 23] == g["kat/identity/ids"][:, 0].tolist()
    c, x = g["kat/duplicates/corpus"], g["kat/duplicates/queries"]
    s, i = oracle.blocked_ip_search(c, x, 6, pad=False)
    rep = oracle.check_topk(g["kat/duplicates/scores"], g["kat/duplicates/ids"], s, i, c, x)
    assert rep.ok, rep.messages
    s, i = oracle.numpy_ip_search(g["kat/toy/corpus"], g["kat/toy/queries"], 10)
    np.testing.assert_array_equal(i, g["kat/toy/ids"])
    assert i.shape == (1, 2)


def test_l2_search_against_bruteforce():
    rng = np.random.default_rng(3)
    c = rng.standard_normal((300, 24)).astype(np.float32)
    x = rng.standard_normal((11, 24)).astype(np.float32)
    s, i = oracle.blocked_l2_search(c, x, 5, block_rows=64)
    d = ((x[:, None, :].astype(np.float64) - c[None].astype(np.float64)) ** 2).sum(-1)
    ref_i = np.argsort(d, axis=1)[:, :5]
    np.testing.assert_array_equal(i, ref_i)
    np.testing.assert_allclose(s, np.take_along_axis(d, ref_i, 1), rtol=1e-4)


def test_check_topk_detects_errors():
    c, x, k, ref_s, ref_i = load_search_case("tiny_ragged")
    bad_i = ref_i.copy()
    worst = int(np.argmin(x[0] @ c.T))
    bad_i[0, 3] = worst
    rep = oracle.check_topk(ref_s, ref_i, ref_s, bad_i, c, x)
    assert not rep.ok
    bad_s = ref_s.copy()
    bad_s[2, 0] *= 1.001
    assert not oracle.check_topk(ref_s, ref_i, bad_s, ref_i, c, x).ok
    assert oracle.check_topk(ref_s, ref_i, ref_s, ref_i, c, x).ok


def test_bf16_rounding_matches_torch():
    import torch
    x = np.random.default_rng(0).standard_normal(10000).astype(np.float32) * 100
    ours = oracle.round_to_bf16(x)
    theirs = torch.from_numpy(x).to(torch.bfloat16).float().numpy()
    np.testing.assert_array_equal(ours, theirs)


# ------------------------------- pooling -------------------------------
POOL_MAP = {"cls": "cls", "mean": "mean", "last": "last", "last_argmax": "last_argmax",
            "mm_cls": "cls", "mm_mean": "mean", "mm_last_token": "last"}


@pytest.mark.parametrize("mask", ["right", "left", "holes", "full"])
@pytest.mark.parametrize("kind", sorted(POOL_MAP))
def test_pooling_matches_reference(kind, mask):
    g = golden()
    h, m = g["pool/hidden"], g[f"pool/mask_{mask}"]
    raw = oracle.pool_and_normalize(h, m, POOL_MAP[kind], normalize=False)
    nrm = oracle.pool_and_normalize(h, m, POOL_MAP[kind], normalize=True)
    np.testing.assert_allclose(raw, g[f"pool/{kind}/{mask}/raw"], rtol=1e-6, atol=1e-7)
    np.testing.assert_allclose(nrm, g[f"pool/{kind}/{mask}/norm"], rtol=1e-6, atol=1e-7)


def test_last_token_variants_diverge_on_interior_zeros():
    g = golden()
    a, b = g["pool/last/holes/raw"], g["pool/last_argmax/holes/raw"]
    assert not np.allclose(a, b)          # decoder/base.py:340 (sum-1) vs mmeb_v2/models.py:148-149 (flip/argmax)
    np.testing.assert_array_equal(g["pool/last/right/raw"], g["pool/last_argmax/right/raw"])


def test_zero_vector_normalise():
    g = golden()
    np.testing.assert_array_equal(oracle.l2_normalize_torch(g["pool/zero_norm_in"]), g["pool/zero_norm_out"])
    assert np.isnan(oracle.l2_normalize_numpy(g["pool/zero_norm_in"])[0]).all()   # base.py:673 has no eps


def test_unknown_pooling_raises():
    with pytest.raises(NotImplementedError):
        oracle.pool_and_normalize(np.zeros((1, 2, 3), np.float32), np.ones((1, 2), np.int64), "max")
