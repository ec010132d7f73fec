"""Helpers for the -m gpu parity tests: every compute call goes through the C ABI (nexus_b200.FlatIndex ->
ctypes -> libnexus_flat.so); the oracle (numpy restatement of the reference) is only the checker."""
import numpy as np

import oracle
from nexus_b200 import FlatIndex
from nexus_b200 import _lib as L

FLAGS = {
    "auto": L.SEARCH_DEFAULT,
    "exact": L.SEARCH_FORCE_EXACT,
    "stream": L.SEARCH_FORCE_STREAM,
    "tensor": L.SEARCH_FORCE_TENSOR,
    "tensor_raw": L.SEARCH_FORCE_TENSOR | L.SEARCH_NO_RESCORE,
    "tensor_1cta": L.SEARCH_FORCE_TENSOR | L.SEARCH_TENSOR_1CTA,
    "tstream": L.SEARCH_FORCE_TSTREAM,
    "tstream_raw": L.SEARCH_FORCE_TSTREAM | L.SEARCH_NO_RESCORE,
}


def stored(c: np.ndarray, store: str) -> np.ndarray:
    """The values the index actually holds (the storage-exact ground truth of SURVEY.md §8(d))."""
    c = c.astype(np.float32)
    if store == "bf16":
        return oracle.round_to_bf16(c)
    if store == "f16":
        return oracle.round_to_f16(c)
    return c


def run_and_check(c, x, k, store="f32", path="auto", rtol=1e-5, atol=1e-6, min_recall=1.0, metric="ip", expect_path=None,
                  add_chunks=1):
    idx = FlatIndex(c.shape[1], metric, store)
    step = (c.shape[0] + add_chunks - 1) // add_chunks if c.shape[0] else 1
    for r0 in range(0, c.shape[0], max(step, 1)):
        idx.add(c[r0:r0 + step])
    assert idx.ntotal == c.shape[0]
    D, I = idx.search(x, k, flags=FLAGS[path])
    assert D.shape == (x.shape[0], k) and I.shape == (x.shape[0], k)
    assert D.dtype == np.float32 and I.dtype == np.int64
    st = idx.stats()
    if expect_path is not None:
        assert st["last_path_name"] == expect_path, st
    cs = stored(c, store)
    xs = x.astype(np.float32)
    if path in ("tensor_raw", "tstream_raw"):
        xs = stored(xs, store)          # the MMA sees the queries rounded to the storage type
        if store == "f32":              # fp32 storage without the re-score: the candidate pass itself, i.e. bf16(q) . bf16(c) (the hi plane)
            cs, xs = oracle.round_to_bf16(cs), oracle.round_to_bf16(xs)
            rtol, atol = max(rtol, 2e-4), max(atol, 2e-5)
    if metric == "ip":
        ref_s, ref_i = oracle.blocked_ip_search(cs, xs, k, pad=True)
        rep = oracle.check_topk(ref_s, ref_i, D, I, cs, xs, rtol=rtol, atol=atol)
    else:
        ref_s, ref_i = oracle.blocked_l2_search(cs, xs, k, pad=True)
        rep = oracle.check_topk(ref_s, ref_i, D, I, None, None, rtol=rtol, atol=atol, tie_tol=1e-4, descending=False)
    assert rep.ok, (st, rep.messages[:8], rep)
    # ids that differ only by a tie within tolerance are not recall losses
    total = max(1, int((I >= 0).sum()))
    assert rep.recall + rep.tie_swaps / total >= min_recall, rep
    return idx, D, I, rep, st
