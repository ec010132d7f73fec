"""MMEB-v2 scoring block on the engine.

Mirrors the ranking section of ``Nexus/evaluation/mmeb_v2/eval_embedding.py:337-404`` (LoRA twin
``eval_emb_lora.py:338``): given the persisted query embeddings ``[Nq, D]``, the candidate dictionary
``{name: [D]}`` and the dataset's ``gt_infos``, build the ``pred_dicts`` that ``RankingMetrics.evaluate`` consumes.

* **global** (``eval_type == "global"``, :351-381): the reference materialises ``scores = q @ cand^T`` ``[Nq, Nc]`` on the GPU,
  fully sorts every row and ships the indices to the host as floats.  Here: one exact flat search for the best ``k``
  candidates per query (``FlatIndex.search``; the ``[Nq, Nc]`` matrix never exists) -- usable at 50 k x 5 M, where the
  reference's form would need 1 TB.
* **local** (else-branch, :382-404): the reference loops over queries, stacks that query's own ``cand_names``, uploads them,
  multiplies and sorts.  Here: the candidate dictionary is uploaded ONCE as an index, the lists become a CSR of row ids and
  one kernel launch ranks every list (``FlatIndex.search_subsets``, kernel KS in ``csrc/search_subset.cu``).

``prediction`` holds the best ``k`` names instead of the complete ranking: every metric reads ``prediction[:k]`` with
``k <= max(k_list)`` = 10 (``utils/eval_utils/metrics.py:12``), so ``k >= 10`` gives identical metrics; pass ``k=None`` for the
complete ranking where it fits (``Nc`` / the longest list <= 2048).  Embeddings are stored as fp32 by default (the pickles
are fp32, ``eval_embedding.py:137``), so scores are fp32 arithmetic like ``torch.matmul``.  No CPU path.
"""
from __future__ import annotations

from typing import Dict, Hashable, List, Optional, Sequence, Tuple

import numpy as np

from . import _lib as L
from .flat import FlatIndex

DEFAULT_K = 10   # max(RankingMetrics.k_list), metrics.py:12


def candidate_matrix(cand_embed_dict: Dict[Hashable, np.ndarray]) -> Tuple[list, np.ndarray]:
    """``cand_keys = list(d.keys())``, ``np.stack([d[key] for key in cand_keys])`` (eval_embedding.py:353-354)."""
    cand_keys = list(cand_embed_dict.keys())
    if not cand_keys:
        raise ValueError("empty candidate dictionary")
    return cand_keys, np.stack([np.asarray(cand_embed_dict[key]) for key in cand_keys])


def build_candidate_index(cand_embed_dict, store_dtype: str = "f32", device=None) -> Tuple[list, Dict[Hashable, int], FlatIndex]:
    """Upload the candidate dictionary once: ``(cand_keys, {name: row}, FlatIndex)``."""
    cand_keys, mat = candidate_matrix(cand_embed_dict)
    idx = FlatIndex(int(mat.shape[-1]), "ip", store_dtype, device=device)
    idx.add(mat)
    return cand_keys, {key: i for i, key in enumerate(cand_keys)}, idx


def rank_global(qry_embeds, cand_embeds, k: Optional[int] = DEFAULT_K, store_dtype: str = "f32", device=None, index: Optional[FlatIndex] = None):
    """-> ``(scores float32 [Nq, k], ranked_idx int64 [Nq, k])`` = the first ``k`` columns of the reference's descending sort of
    ``q @ cand^T`` (eval_embedding.py:358-362), -1 padded when ``k > Nc``."""
    if index is None:
        cand_embeds = np.asarray(cand_embeds) if not hasattr(cand_embeds, "is_cuda") else cand_embeds
        index = FlatIndex(int(cand_embeds.shape[-1]), "ip", store_dtype, device=device)
        index.add(cand_embeds)
    nc = index.ntotal
    if k is None:
        k = nc
    if k > L.MAX_K:
        raise ValueError(f"k={k}: at most {L.MAX_K} ranked candidates per query (the metrics read the first {DEFAULT_K})")
    return index.search(qry_embeds, int(k))


def lists_to_csr(cand_lists: Sequence[Sequence[Hashable]], row_of: Dict[Hashable, int]) -> Tuple[np.ndarray, np.ndarray]:
    """Per-query name lists -> CSR ``(offsets int64 [Nq + 1], rows int64 [total])``.  An unknown name raises ``KeyError``
    exactly where the reference does (``cand_embed_dict[name]``, eval_embedding.py:390)."""
    offsets = np.zeros(len(cand_lists) + 1, dtype=np.int64)
    for i, names in enumerate(cand_lists):
        offsets[i + 1] = offsets[i] + len(names)
    rows = np.empty(int(offsets[-1]), dtype=np.int64)
    p = 0
    for names in cand_lists:
        for name in names:
            rows[p] = row_of[name]
            p += 1
    return offsets, rows


def rank_local(qry_embeds, cand_lists: Sequence[Sequence[Hashable]], row_of: Dict[Hashable, int], index: FlatIndex,
               k: Optional[int] = DEFAULT_K):
    """-> ``(scores float32 [Nq, k], ranked_idx int64 [Nq, k])``: per query the first ``k`` positions of the descending sort of
    its own candidates' scores (eval_embedding.py:393-396), -1 padded beyond the list length."""
    if len(cand_lists) != len(qry_embeds):
        raise ValueError(f"{len(cand_lists)} candidate lists for {len(qry_embeds)} queries")
    offsets, rows = lists_to_csr(cand_lists, row_of)
    if k is None:
        k = int(np.diff(offsets).max()) if len(cand_lists) else 1
    k = max(int(k), 1)
    if k > L.MAX_K:
        raise ValueError(f"k={k}: at most {L.MAX_K} ranked candidates per query (the metrics read the first {DEFAULT_K})")
    return index.search_subsets(qry_embeds, offsets, rows, k)


def _labels(gt_info):
    rel_docids = gt_info["label_name"] if isinstance(gt_info["label_name"], list) else [gt_info["label_name"]]   # :371,398
    return rel_docids, gt_info.get("rel_scores", None)                                                            # :372,399


def build_pred_dicts(qry_embeds, cand_embed_dict, gt_infos: List[dict], eval_type: str = "global", k: Optional[int] = DEFAULT_K,
                     store_dtype: str = "f32", device=None) -> List[dict]:
    """The ``pred_dicts`` of eval_embedding.py:345-404 with ``prediction`` = the best ``k`` candidate names."""
    qry_embeds = np.asarray(qry_embeds)
    if len(gt_infos) != len(qry_embeds):
        raise ValueError(f"{len(gt_infos)} gt_infos for {len(qry_embeds)} queries")
    cand_keys, row_of, index = build_candidate_index(cand_embed_dict, store_dtype, device)
    pred_dicts = []
    if eval_type == "global":                                                   # :350
        _, ranked = rank_global(qry_embeds, None, k, index=index)
        for ranked_idx, gt_info in zip(ranked, gt_infos):
            rel_docids, rel_scores = _labels(gt_info)
            pred_dicts.append({"prediction": [cand_keys[int(i)] for i in ranked_idx if i >= 0],      # :375
                               "label": rel_docids, "rel_scores": rel_scores})
    else:
        lists = [gt_info["cand_names"] for gt_info in gt_infos]                # :389
        _, ranked = rank_local(qry_embeds, lists, row_of, index, k)
        for ranked_idx, names, gt_info in zip(ranked, lists, gt_infos):
            rel_docids, rel_scores = _labels(gt_info)
            pred_dicts.append({"prediction": [names[int(i)] for i in ranked_idx if i >= 0],          # :402
                               "label": rel_docids, "rel_scores": rel_scores})
    return pred_dicts
