"""CPU restatement of the reference's flat inner-product search.  TEST INFRASTRUCTURE ONLY
(see ``oracle/__init__.py`` for who may import this and for the parity status).

Each function cites the reference lines (under ``/root/reference``) it follows.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Callable, List, Optional, Tuple

import numpy as np

FLT_MAX = float(np.finfo(np.float32).max)


# --------------------------------------------------------------------------------------
# The reference's own implementation
# --------------------------------------------------------------------------------------
def numpy_ip_search(embeddings: np.ndarray, queries: np.ndarray, k: int) -> Tuple[np.ndarray, np.ndarray]:
    """Restates ``NumpyIPIndex.__init__`` + ``.search``
    (Nexus/evaluation/text_retrieval/utils.py:16-31).

    fp32 cast of both sides (:18, :21), ``scores = q @ emb.T`` (:22), ``top_k = min(k, N)`` (:23),
    empty arrays when ``top_k <= 0`` (:24-27), full descending argsort and column cut (:29),
    ``take_along_axis`` (:30).  Note the columns are TRUNCATED to ``min(k, N)``, not padded.
    """
    emb = embeddings.astype(np.float32)
    q = queries.astype(np.float32)
    scores = q @ emb.T
    top_k = min(k, emb.shape[0])
    if top_k <= 0:
        return (np.empty((q.shape[0], 0), dtype=np.float32), np.empty((q.shape[0], 0), dtype=np.int64))
    order = np.argsort(-scores, axis=1)[:, :top_k]
    return np.take_along_axis(scores, order, axis=1), order


def reference_chunked_search(index_search: Callable[[np.ndarray, int], Tuple[np.ndarray, np.ndarray]],
                             queries: np.ndarray, k: int, chunk: int = 32) -> Tuple[np.ndarray, np.ndarray]:
    """Restates ``search()`` (utils.py:194-230): 32-row query chunks (:221-223), each cast to fp32
    and sent to ``faiss_index.search`` (:224), results concatenated along axis 0 (:228-229)."""
    all_s: List[np.ndarray] = []
    all_i: List[np.ndarray] = []
    for i in range(0, len(queries), chunk):
        s, ix = index_search(queries[i:i + chunk].astype(np.float32), k)
        all_s.append(s)
        all_i.append(ix)
    return np.concatenate(all_s, axis=0), np.concatenate(all_i, axis=0)


# --------------------------------------------------------------------------------------
# Faiss conventions (third-party; restated from its published behaviour)
# --------------------------------------------------------------------------------------
def faiss_pad(scores: np.ndarray, ids: np.ndarray, k: int, metric: str = "ip") -> Tuple[np.ndarray, np.ndarray]:
    """Faiss ``IndexFlat.search`` always returns ``[nq, k]``: when ``k > ntotal`` the tail is filled
    with id ``-1`` and the heap's neutral score (``-FLT_MAX`` for inner product, ``+FLT_MAX`` for L2).
    The reference tolerates that padding at searcher.py:156 (``if indice != -1``)."""
    nq, have = scores.shape
    if have >= k:
        return scores[:, :k], ids[:, :k]
    fill = -FLT_MAX if metric == "ip" else FLT_MAX
    s = np.full((nq, k), fill, dtype=np.float32)
    i = np.full((nq, k), -1, dtype=np.int64)
    s[:, :have] = scores
    i[:, :have] = ids
    return s, i


def _merge_desc(best_s, best_i, blk_s, blk_i, k):
    s = np.concatenate([best_s, blk_s], axis=1)
    i = np.concatenate([best_i, blk_i], axis=1)
    if s.shape[1] > k:
        part = np.argpartition(-s, k - 1, axis=1)[:, :k]
        s = np.take_along_axis(s, part, axis=1)
        i = np.take_along_axis(i, part, axis=1)
    return s, i


def blocked_ip_search(embeddings: np.ndarray, queries: np.ndarray, k: int, block_rows: int = 65536,
                      pad: bool = True, q_block: int = 1024) -> Tuple[np.ndarray, np.ndarray]:
    """Same arithmetic as ``NumpyIPIndex.search`` (utils.py:18,21-23,29-30) evaluated in corpus row
    blocks so the ``[Q,N]`` fp32 score matrix is never materialised: per block ``q @ blk.T`` in fp32,
    ``argpartition`` to the block's top-k, merge with the running top-k, final descending sort
    (ties: lower id first, which is one of the orders the reference's non-stable argsort may give).
    ``pad=True`` returns Faiss-shaped ``[Q,k]`` with ``-1`` padding; ``pad=False`` truncates like
    the reference's numpy fallback (utils.py:23)."""
    emb = embeddings
    n = emb.shape[0]
    nq = queries.shape[0]
    kk = min(k, n)
    if kk <= 0:
        s = np.empty((nq, 0), dtype=np.float32)
        i = np.empty((nq, 0), dtype=np.int64)
        return faiss_pad(s, i, k) if pad and k > 0 else (s, i)
    out_s = np.empty((nq, kk), dtype=np.float32)
    out_i = np.empty((nq, kk), dtype=np.int64)
    for q0 in range(0, nq, q_block):
        q = queries[q0:q0 + q_block].astype(np.float32)
        best_s = np.empty((q.shape[0], 0), dtype=np.float32)
        best_i = np.empty((q.shape[0], 0), dtype=np.int64)
        for r0 in range(0, n, block_rows):
            blk = emb[r0:r0 + block_rows].astype(np.float32, copy=False)   # utils.py:18 casts once, up front
            sc = q @ blk.T
            kb = min(kk, sc.shape[1])
            if sc.shape[1] > kb:
                part = np.argpartition(-sc, kb - 1, axis=1)[:, :kb]
                bs = np.take_along_axis(sc, part, axis=1)
                bi = part.astype(np.int64) + r0
            else:
                bs = sc
                bi = np.broadcast_to(np.arange(r0, r0 + sc.shape[1], dtype=np.int64), sc.shape).copy()
            best_s, best_i = _merge_desc(best_s, best_i, bs, bi, kk)
        # final order: score descending, id ascending among equal scores
        order = np.lexsort((best_i, -best_s), axis=1)
        out_s[q0:q0 + q.shape[0]] = np.take_along_axis(best_s, order, axis=1)
        out_i[q0:q0 + q.shape[0]] = np.take_along_axis(best_i, order, axis=1)
    return faiss_pad(out_s, out_i, k) if pad else (out_s, out_i)


def blocked_ip_search_thr(embeddings: np.ndarray, queries: np.ndarray, k: int, block_rows: int = 65536,
                          pad: bool = True, q_block: int = 1024) -> Tuple[np.ndarray, np.ndarray]:
    """``blocked_ip_search`` with the selection a CPU flat index (Faiss's sgemm + per-row heap) would use: the same fp32
    ``q @ blk.T`` per corpus block (utils.py:21-22), but from the second block on only scores that beat the row's running
    k-th best are touched (one vectorised compare per block instead of an ``argpartition`` of every block), and the running
    lists are cut back to k only when they have doubled.  Same results as ``blocked_ip_search`` (same products, same final
    order; asserted in tests/test_oracle_properties_cpu.py); it exists so that the timed CPU baseline is bound by the sgemm,
    as real Faiss is, and not by numpy's selection."""
    emb = embeddings
    n = emb.shape[0]
    nq = queries.shape[0]
    kk = min(k, n)
    if kk <= 0:
        return blocked_ip_search(embeddings, queries, k, block_rows, pad, q_block)
    out_s = np.empty((nq, kk), dtype=np.float32)
    out_i = np.empty((nq, kk), dtype=np.int64)
    for q0 in range(0, nq, q_block):
        q = queries[q0:q0 + q_block].astype(np.float32)
        m = q.shape[0]
        cap = 2 * kk
        best_s = np.full((m, cap), -np.inf, dtype=np.float32)
        best_i = np.full((m, cap), -1, dtype=np.int64)
        cnt = np.zeros(m, dtype=np.int64)
        thr = np.full(m, -np.inf, dtype=np.float32)          # the row's k-th best so far (a lower bound of the final one)
        for r0 in range(0, n, block_rows):
            blk = emb[r0:r0 + block_rows].astype(np.float32, copy=False)   # utils.py:18 casts once, up front
            sc = q @ blk.T
            rows, cols = np.nonzero(sc > thr[:, None]) if r0 else (None, None)
            if r0 == 0:
                kb = min(kk, sc.shape[1])
                part = np.argpartition(-sc, kb - 1, axis=1)[:, :kb] if sc.shape[1] > kb else np.broadcast_to(np.arange(kb), (m, kb))
                best_s[:, :kb] = np.take_along_axis(sc, part, axis=1)
                best_i[:, :kb] = part
                cnt[:] = kb
                if kb == kk:
                    thr = best_s[:, :kk].min(axis=1)
                continue
            if rows.size == 0:
                continue
            vals = sc[rows, cols]
            # survivors arrive row-major: append each row's run behind what it holds, pruning rows that would overflow
            starts = np.flatnonzero(np.r_[True, rows[1:] != rows[:-1]])
            ends = np.r_[starts[1:], rows.size]
            for a, b in zip(starts, ends):
                r = rows[a]
                take = b - a
                if cnt[r] + take > cap:
                    ss = np.concatenate([best_s[r, :cnt[r]], vals[a:b]])
                    ii = np.concatenate([best_i[r, :cnt[r]], cols[a:b] + r0])
                    keep = np.argpartition(-ss, kk - 1)[:kk]
                    best_s[r, :kk], best_i[r, :kk] = ss[keep], ii[keep]
                    best_s[r, kk:], best_i[r, kk:] = -np.inf, -1
                    cnt[r] = kk
                else:
                    best_s[r, cnt[r]:cnt[r] + take] = vals[a:b]
                    best_i[r, cnt[r]:cnt[r] + take] = cols[a:b] + r0
                    cnt[r] += take
                if cnt[r] >= kk:
                    thr[r] = np.partition(best_s[r, :cnt[r]], cnt[r] - kk)[cnt[r] - kk]
        order = np.lexsort((best_i, -best_s), axis=1)[:, :kk]
        out_s[q0:q0 + m] = np.take_along_axis(best_s, order, axis=1)
        out_i[q0:q0 + m] = np.take_along_axis(best_i, order, axis=1)
    return faiss_pad(out_s, out_i, k) if pad else (out_s, out_i)


def blocked_l2_search(embeddings: np.ndarray, queries: np.ndarray, k: int, block_rows: int = 65536,
                      pad: bool = True) -> Tuple[np.ndarray, np.ndarray]:
    """Faiss ``IndexFlatL2`` semantics, used by ``BaseRetriever.eval_step`` for ``EuclideanScorer``
    (Nexus/training/embedder/recommendation/modeling.py:185-186,197): squared Euclidean distance,
    rows sorted ASCENDING.  Computed as ``|q|^2 - 2 q.c + |c|^2`` in fp32 like Faiss' BLAS path."""
    n = embeddings.shape[0]
    nq = queries.shape[0]
    kk = min(k, n)
    q = queries.astype(np.float32)
    qn = (q * q).sum(axis=1, keepdims=True)
    best_s = np.empty((nq, 0), dtype=np.float32)
    best_i = np.empty((nq, 0), dtype=np.int64)
    for r0 in range(0, n, block_rows):
        blk = embeddings[r0:r0 + block_rows].astype(np.float32)
        d = qn - 2.0 * (q @ blk.T) + (blk * blk).sum(axis=1)[None, :]
        bi = np.broadcast_to(np.arange(r0, r0 + d.shape[1], dtype=np.int64), d.shape)
        s = np.concatenate([best_s, d.astype(np.float32)], axis=1)
        i = np.concatenate([best_i, bi], axis=1)
        if s.shape[1] > kk:
            part = np.argpartition(s, kk - 1, axis=1)[:, :kk]
            s = np.take_along_axis(s, part, axis=1)
            i = np.take_along_axis(i, part, axis=1)
        best_s, best_i = s, i
    order = np.lexsort((best_i, best_s), axis=1)
    best_s = np.take_along_axis(best_s, order, axis=1)
    best_i = np.take_along_axis(best_i, order, axis=1)
    return faiss_pad(best_s, best_i, k, metric="l2") if pad else (best_s, best_i)


# --------------------------------------------------------------------------------------
# float64 adjudicator: decides which disagreements are "ties within tolerance"
# --------------------------------------------------------------------------------------
def exact_scores_f64(embeddings: np.ndarray, queries: np.ndarray, ids: np.ndarray) -> np.ndarray:
    """float64 inner products of query r with rows ``ids[r, :]`` (id -1 -> -inf)."""
    nq, k = ids.shape
    out = np.full((nq, k), -np.inf, dtype=np.float64)
    q64 = queries.astype(np.float64)
    for r in range(nq):
        valid = ids[r] >= 0
        if valid.any():
            rows = embeddings[ids[r][valid]].astype(np.float64)
            out[r, valid] = rows @ q64[r]
    return out


@dataclass
class TopkReport:
    ok: bool
    n_rows: int
    exact_id_rows: int            # rows whose id list equals the reference's position by position
    set_equal_rows: int           # rows whose id SET equals the reference's
    tie_swaps: int                # ids that differ but are justified by a tie within tolerance
    bad_ids: int                  # ids that differ and are NOT justified
    max_score_rel_err: float
    unsorted_rows: int
    recall: float                 # |got ∩ ref| / |ref| over valid ids
    messages: List[str] = field(default_factory=list)

    def __bool__(self) -> bool:  # pragma: no cover - convenience
        return self.ok


def check_topk(ref_scores: np.ndarray, ref_ids: np.ndarray, got_scores: np.ndarray, got_ids: np.ndarray,
               embeddings: Optional[np.ndarray] = None, queries: Optional[np.ndarray] = None,
               rtol: float = 1e-5, atol: float = 1e-6, tie_tol: Optional[float] = None,
               descending: bool = True) -> TopkReport:
    """Tie-aware comparison of a top-k result with the oracle's.

    The bar (BASELINE.json north_star): ids bit-exact except among scores tied within tolerance,
    fp32 scores within ``rtol`` relative.  Neither the reference's ``np.argsort`` (non-stable
    quicksort, utils.py:29) nor Faiss' heap defines an order among equal scores, so a differing id
    is accepted iff its float64 score (recomputed from ``embeddings``/``queries`` when given, else
    the reported score) is within ``tie_tol`` of the reference's k-th score for that row.
    """
    assert ref_ids.shape == got_ids.shape, (ref_ids.shape, got_ids.shape)
    assert ref_scores.shape == got_scores.shape
    nq, k = ref_ids.shape
    msgs: List[str] = []
    if tie_tol is None:
        tie_tol = 4 * atol
    sign = 1.0 if descending else -1.0
    exact_rows = int((ref_ids == got_ids).all(axis=1).sum()) if k else nq
    set_rows = 0
    tie_swaps = 0
    bad = 0
    unsorted_rows = 0
    inter = 0
    total = 0
    max_rel = 0.0
    have_data = embeddings is not None and queries is not None and descending
    got64 = exact_scores_f64(embeddings, queries, got_ids) if have_data else got_scores.astype(np.float64)
    for r in range(nq):
        gs = got_scores[r].astype(np.float64)
        gi = got_ids[r]
        rs = ref_scores[r].astype(np.float64)
        ri = ref_ids[r]
        valid_g = gi >= 0
        valid_r = ri >= 0
        if valid_g.sum() != valid_r.sum():
            bad += abs(int(valid_g.sum()) - int(valid_r.sum()))
            msgs.append(f"row {r}: {int(valid_g.sum())} valid ids, reference has {int(valid_r.sum())}")
        # padding must trail
        if valid_g.any() and not valid_g[: int(valid_g.sum())].all():
            bad += 1
            msgs.append(f"row {r}: -1 padding is not trailing")
        v = gs[valid_g] * sign
        if v.size > 1 and (np.diff(v) > atol + rtol * np.abs(v[1:])).any():
            unsorted_rows += 1
        # reported scores must be the true scores of the reported ids
        if have_data and valid_g.any():
            err = np.abs(gs[valid_g] - got64[r][valid_g])
            rel = err / np.maximum(np.abs(got64[r][valid_g]), atol / max(rtol, 1e-30))
            max_rel = max(max_rel, float(rel.max()))
        elif valid_g.any() and valid_r.any():
            m = min(int(valid_g.sum()), int(valid_r.sum()))
            err = np.abs(gs[:m] - rs[:m])
            rel = err / np.maximum(np.abs(rs[:m]), atol / max(rtol, 1e-30))
            max_rel = max(max_rel, float(rel.max()))
        sg = set(gi[valid_g].tolist())
        sr = set(ri[valid_r].tolist())
        if len(sg) != int(valid_g.sum()):
            bad += 1
            msgs.append(f"row {r}: duplicate ids in result")
        inter += len(sg & sr)
        total += len(sr)
        if sg == sr:
            set_rows += 1
            continue
        if not valid_r.any():
            continue
        kth = rs[valid_r][-1] * sign           # the reference's worst kept score
        pos = {int(i): j for j, i in enumerate(gi.tolist())}
        for e in sg - sr:
            sc = got64[r][pos[e]] * sign
            if sc >= kth - tie_tol - rtol * abs(kth):
                tie_swaps += 1
            else:
                bad += 1
                if len(msgs) < 20:
                    msgs.append(f"row {r}: id {e} score {sc:.9g} below reference k-th {kth:.9g}")
    ok = bad == 0 and unsorted_rows == 0 and max_rel <= rtol
    if max_rel > rtol:
        msgs.append(f"max relative score error {max_rel:.3e} > rtol {rtol:.1e}")
    if unsorted_rows:
        msgs.append(f"{unsorted_rows} rows not sorted")
    return TopkReport(ok=ok, n_rows=nq, exact_id_rows=exact_rows, set_equal_rows=set_rows,
                      tie_swaps=tie_swaps, bad_ids=bad, max_score_rel_err=max_rel,
                      unsorted_rows=unsorted_rows, recall=(inter / total if total else 1.0), messages=msgs)


# --------------------------------------------------------------------------------------
# Storage rounding + synthetic data
# --------------------------------------------------------------------------------------
def round_to_bf16(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 (round to nearest even) -> fp32, bit-exact with CUDA's __float2bfloat16_rn."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32)
    nan = (u & 0x7FFFFFFF) > 0x7F800000
    r = (u + (0x7FFF + ((u >> 16) & 1))) & 0xFFFF0000
    r = np.where(nan, u | 0x00400000, r).astype(np.uint32)
    return r.view(np.float32)


def round_to_f16(x: np.ndarray) -> np.ndarray:
    return np.asarray(x, dtype=np.float32).astype(np.float16).astype(np.float32)


def make_unit_gaussian(n: int, d: int, seed: int, normalize: bool = True, chunk: int = 1 << 18) -> np.ndarray:
    """Isotropic N(0,1) rows, optionally L2-normalised (configs C1-C4); C5 passes normalize=False."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, d), dtype=np.float32)
    for r0 in range(0, n, chunk):
        x = rng.standard_normal((min(chunk, n - r0), d), dtype=np.float32)
        if normalize:
            x /= np.linalg.norm(x, axis=1, keepdims=True)
        out[r0:r0 + x.shape[0]] = x
    return out


def make_clustered(n: int, d: int, seed: int, n_centroids: int = 4096, noise: float = 0.5,
                   centroid_seed: int = 7) -> np.ndarray:
    """Clustered variant: unit centroids + N(0, noise^2/d) noise, renormalised."""
    crng = np.random.default_rng(centroid_seed)
    cent = crng.standard_normal((n_centroids, d), dtype=np.float32)
    cent /= np.linalg.norm(cent, axis=1, keepdims=True)
    rng = np.random.default_rng(seed)
    which = rng.integers(0, n_centroids, size=n)
    x = cent[which] + rng.standard_normal((n, d), dtype=np.float32) * np.float32(noise / np.sqrt(d))
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    return x.astype(np.float32)
