"""CPU restatements of the two callers right after the search.  TEST INFRASTRUCTURE ONLY.

* ``unpack_results_loop``  = the nested loop at Nexus/evaluation/text_retrieval/searcher.py:152-159, statement for statement
* ``positive_pred``        = ``torch.cat([pos, topk])`` + stable descending ``torch.sort`` + ``indice[:, :k] == 0``
                             at Nexus/training/embedder/recommendation/modeling.py:199-205, in numpy
                             (``np.argsort(kind="stable")`` on the negated scores is the stable descending order)
Parity: both are pinned by construction against torch / plain Python semantics (the torch form is cross-checked in
tests/test_callers_cpu.py with the reference's own three lines).
"""
from __future__ import annotations

import numpy as np


def unpack_results_loop(queries_ids, corpus_ids, all_scores, all_indices, ignore_identical_ids=False):
    results = {}
    for idx, (scores, indices) in enumerate(zip(all_scores, all_indices)):          # searcher.py:152
        results[queries_ids[idx]] = {}                                             # :153
        for score, indice in zip(scores, indices):                                 # :154
            if indice != -1:                                                       # :155
                if ignore_identical_ids and corpus_ids[indice] == queries_ids[idx]:  # :156
                    continue
                results[queries_ids[idx]][corpus_ids[indice]] = float(score)       # :158
    return results


def positive_pred(pos_scores: np.ndarray, topk_scores: np.ndarray):
    """-> (rank int32 [B], pred bool [B, k]) with pred = (indice[:, :k] == 0) of the reference."""
    pos = np.asarray(pos_scores, dtype=np.float32).reshape(-1, 1)                   # modeling.py:199-200
    topk = np.asarray(topk_scores, dtype=np.float32)
    all_scores = np.concatenate([pos, topk], axis=-1)                              # :201
    indice = np.argsort(-all_scores, axis=-1, kind="stable")                       # :203 (stable, descending)
    k = topk.shape[1]
    pred = indice[:, :k] == 0                                                      # :204
    rank = np.argmax(indice == 0, axis=-1).astype(np.int32)                        # position of the positive, in [0, k]
    return rank, pred
