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


# ---- MMEB-v2 ranking block (Nexus/evaluation/mmeb_v2/eval_embedding.py:345-404) -------------------------------------------
# numpy restatement; `compute_similarity` is torch.matmul(q, p.T) (models.py:186-188).  torch.sort(descending=True) is not
# stable, so exactly tied scores may come out in any order in the reference; the restatement uses the stable order
# (ascending position among ties), which is one of the orders the reference may produce.  Pinned against the reference's own
# lines run here (tests/golden/make_golden_mmeb.py -> tests/golden/mmeb_golden.npz).

def mmeb_similarity(q: np.ndarray, p: np.ndarray) -> np.ndarray:
    return np.asarray(q, dtype=np.float32) @ np.asarray(p, dtype=np.float32).T      # models.py:188


def mmeb_global_ranking(qry_embeds, cand_embed_dict):
    """-> (cand_keys, scores [Nq, Nc], ranked_indices [Nq, Nc]) as at eval_embedding.py:353-362."""
    cand_keys = list(cand_embed_dict.keys())                                       # :353
    cand_embeds = np.stack([cand_embed_dict[key] for key in cand_keys])            # :354
    scores = mmeb_similarity(qry_embeds, cand_embeds)                              # :359
    ranked_indices = np.argsort(-scores, axis=1, kind="stable")                    # :361
    return cand_keys, scores, ranked_indices


def mmeb_local_ranking(qry_embeds, cand_embed_dict, gt_infos):
    """-> list of (sim_scores [Nc_i], ranked_idx [Nc_i]) per query as at eval_embedding.py:389-396."""
    out = []
    for qry_vec, gt_info in zip(np.asarray(qry_embeds), gt_infos):                 # :384-385
        cand_names = gt_info["cand_names"]                                         # :389
        cand_embeds = np.stack([cand_embed_dict[name] for name in cand_names])     # :390
        sim_scores = mmeb_similarity(qry_vec[None, :], cand_embeds)[0]             # :394
        ranked_idx = np.argsort(-sim_scores, kind="stable")                        # :395
        out.append((sim_scores, ranked_idx))
    return out


def mmeb_pred_dicts(qry_embeds, cand_embed_dict, gt_infos, eval_type="global"):
    """The complete-ranking ``pred_dicts`` of eval_embedding.py:345-404."""
    pred_dicts = []
    if eval_type == "global":                                                      # :349-350
        cand_keys, _, ranked_indices = mmeb_global_ranking(qry_embeds, cand_embed_dict)
        for ranked_idx, gt_info in zip(ranked_indices, gt_infos):                  # :367
            rel_docids = gt_info["label_name"] if isinstance(gt_info["label_name"], list) else [gt_info["label_name"]]   # :371
            pred_dicts.append({"prediction": [cand_keys[int(i)] for i in ranked_idx],                                    # :375
                               "label": rel_docids, "rel_scores": gt_info.get("rel_scores", None)})
    else:
        for (_, ranked_idx), gt_info in zip(mmeb_local_ranking(qry_embeds, cand_embed_dict, gt_infos), gt_infos):
            cand_names = gt_info["cand_names"]
            rel_docids = gt_info["label_name"] if isinstance(gt_info["label_name"], list) else [gt_info["label_name"]]   # :398
            pred_dicts.append({"prediction": [cand_names[int(i)] for i in ranked_idx],                                   # :402
                               "label": rel_docids, "rel_scores": gt_info.get("rel_scores", None)})
    return pred_dicts
