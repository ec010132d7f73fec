"""CPU oracle for the flat-search hot path.  TEST INFRASTRUCTURE ONLY.

Nothing under ``nexus_b200/`` may import this package.  The only legal
importers are ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s
``cpu_baseline`` / ``--impl reference`` legs, and there only as the checker or
the timed CPU baseline -- never as the product path.

Parity status: PINNED against outputs of the reference's own code
(``Nexus/evaluation/text_retrieval/utils.py`` ``NumpyIPIndex``/``index``/``search``
and the embedders' ``pooling`` + ``F.normalize``) run in the build container;
the vectors and the generating script live in ``tests/golden/``.
The Faiss arithmetic the reference prefers when ``faiss`` is importable is a
third-party dependency (``faiss-cpu``, unpinned in ``setup.py:30``; ``==1.13.2`` in
``examples/multimodal_retrieval/requirements.txt:38``) that is absent here; its
published flat-index semantics (exact inner product / squared L2, rows sorted,
``-1`` id padding when ``k > ntotal``) are restated in ``flat_search.py`` and that
part is "parity unpinned" against Faiss itself.
"""
from .flat_search import (  # noqa: F401
    numpy_ip_search,
    reference_chunked_search,
    blocked_ip_search,
    blocked_ip_search_thr,
    blocked_l2_search,
    faiss_pad,
    exact_scores_f64,
    check_topk,
    TopkReport,
    round_to_bf16,
    round_to_f16,
    make_unit_gaussian,
    make_clustered,
)
from .pooling import (  # noqa: F401
    pool_cls,
    pool_mean,
    pool_last_token,
    pool_last_token_argmax,
    l2_normalize_torch,
    l2_normalize_numpy,
    pool_and_normalize,
)
from .callers import (  # noqa: F401
    unpack_results_loop,
    positive_pred,
    mmeb_similarity,
    mmeb_global_ranking,
    mmeb_local_ranking,
    mmeb_pred_dicts,
)
