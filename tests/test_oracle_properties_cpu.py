"""Property tests of the oracle itself (CPU, hypothesis): the blocked restatement used at sizes where [Q, N] does not fit must agree
with the direct restatement of NumpyIPIndex whatever the block sizes, and the MMEB local ranking must be the global ranking
restricted to the query's own list.  Continuous random inputs: no exact ties, so ids are compared exactly."""
import numpy as np
import pytest
from hypothesis import given, settings, strategies as st

import oracle


@settings(max_examples=40, deadline=None)
@given(n=st.integers(1, 400), d=st.sampled_from([1, 7, 32, 100]), q=st.integers(1, 40), k=st.integers(1, 60),
       block_rows=st.sampled_from([1, 13, 64, 65536]), q_block=st.sampled_from([1, 5, 1024]), seed=st.integers(0, 2**31 - 1))
def test_blocked_search_equals_direct_search(n, d, q, k, block_rows, q_block, seed):
    rng = np.random.default_rng(seed)
    c = rng.standard_normal((n, d)).astype(np.float32)
    x = rng.standard_normal((q, d)).astype(np.float32)
    ds, di = oracle.numpy_ip_search(c, x, k)                      # columns truncated to min(k, N), utils.py:23
    bs, bi = oracle.blocked_ip_search(c, x, k, block_rows=block_rows, q_block=q_block, pad=True)
    assert bs.shape == (q, k) and bi.shape == (q, k)
    m = min(k, n)
    np.testing.assert_allclose(bs[:, :m], ds, rtol=1e-5, atol=1e-5)
    full = x @ c.T
    for r in range(q):                                           # same ids unless two scores are closer than the summation noise
        if not np.array_equal(bi[r, :m], di[r]):
            np.testing.assert_allclose(full[r, bi[r, :m]], full[r, di[r]], rtol=1e-5, atol=1e-5)
    assert np.all(bi[:, m:] == -1) and np.all(bs[:, m:] == -np.finfo(np.float32).max)   # Faiss padding
    ts, ti = oracle.blocked_ip_search(c, x, k, block_rows=block_rows, q_block=q_block, pad=False)
    assert ts.shape == (q, m) and np.array_equal(ti, bi[:, :m])


@settings(max_examples=25, deadline=None)
@given(nc=st.integers(2, 120), d=st.sampled_from([4, 48]), nq=st.integers(1, 12), seed=st.integers(0, 2**31 - 1))
def test_mmeb_local_ranking_is_the_global_ranking_restricted_to_the_list(nc, d, nq, seed):
    rng = np.random.default_rng(seed)
    cand = rng.standard_normal((nc, d)).astype(np.float32)
    qry = rng.standard_normal((nq, d)).astype(np.float32)
    names = [f"c{i}" for i in range(nc)]
    cdict = {n: cand[i] for i, n in enumerate(names)}
    gts = []
    for _ in range(nq):
        pick = rng.choice(nc, size=int(rng.integers(1, nc + 1)), replace=False)
        gts.append({"label_name": names[int(pick[0])], "cand_names": [names[int(j)] for j in pick]})
    keys, scores, ranked = oracle.mmeb_global_ranking(qry, cdict)
    local = oracle.mmeb_local_ranking(qry, cdict, gts)
    for qi, ((sc, rk), gt) in enumerate(zip(local, gts)):
        listed = set(gt["cand_names"])
        expect = [keys[j] for j in ranked[qi] if keys[j] in listed]
        got = [gt["cand_names"][j] for j in rk]
        if got != expect:                                        # only summation noise between the [1, Nc_i] and [Nq, Nc] products may differ
            a = np.array([scores[qi][keys.index(n)] for n in got])
            b = np.array([scores[qi][keys.index(n)] for n in expect])
            np.testing.assert_allclose(a, b, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("n,d,q,k,block", [(5000, 32, 50, 10, 512), (70000, 48, 33, 100, 8192), (300, 16, 7, 500, 64), (20000, 24, 9, 1, 4096)])
def test_threshold_filtered_search_equals_blocked_search(n, d, q, k, block):
    """The timed CPU baseline (bench.py) uses blocked_ip_search_thr: same products, same results as blocked_ip_search."""
    c = oracle.make_unit_gaussian(n, d, seed=n % 97)
    x = oracle.make_unit_gaussian(q, d, seed=q)
    a = oracle.blocked_ip_search(c, x, k, block_rows=block)
    b = oracle.blocked_ip_search_thr(c, x, k, block_rows=block)
    np.testing.assert_array_equal(a[1], b[1])
    np.testing.assert_array_equal(a[0], b[0])
