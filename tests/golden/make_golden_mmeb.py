"""Golden vectors for the MMEB-v2 ranking block, produced by the REFERENCE's own code in the build container.

    python tests/golden/make_golden_mmeb.py        ->  tests/golden/mmeb_golden.npz

Executed from ``/root/reference`` (read-only, absent on the GPU box -- hence the committed fixture):
  * ``MMEBEmbeddingModel.compute_similarity`` (Nexus/evaluation/mmeb_v2/models.py:186-188), called unbound;
  * ``RankingMetrics.evaluate`` (Nexus/evaluation/mmeb_v2/utils/eval_utils/metrics.py:191-266), unmodified.
The ranking statements themselves sit inside ``main()`` of eval_embedding.py (:345-404, not callable on their own); they are
replayed here statement for statement with the reference's functions: ``np.stack`` of the dictionary, ``torch.from_numpy``,
``compute_similarity``, ``torch.sort(descending=True)``, ``.cpu().float().numpy()``, the ``pred_dicts`` comprehension.
Inputs come from tests/_common.py::gen_mmeb_case (seeded), their sha256 is stored beside the outputs.
"""
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, "/root/reference")

from _common import MMEB_CASES, gen_mmeb_case, sha  # noqa: E402
from Nexus.evaluation.mmeb_v2.models import MMEBEmbeddingModel  # noqa: E402
from Nexus.evaluation.mmeb_v2.utils.eval_utils.metrics import RankingMetrics  # noqa: E402

METRICS = ["hit", "ndcg", "precision", "recall", "f1", "map", "mrr"]   # eval_embedding.py:407 default


def compute_similarity(q, p):
    return MMEBEmbeddingModel.compute_similarity(None, q, p)


def run_global(qry_embeds, cand_embed_dict, gt_infos):
    qry_tensor = torch.from_numpy(qry_embeds)                                           # :346
    cand_keys = list(cand_embed_dict.keys())                                            # :353
    cand_embeds = np.stack([cand_embed_dict[key] for key in cand_keys])                 # :354
    cand_tensor = torch.from_numpy(cand_embeds)                                         # :355
    with torch.no_grad():
        scores = compute_similarity(qry_tensor, cand_tensor)                            # :359
        _, ranked_indices = torch.sort(scores, dim=1, descending=True)                  # :361
        ranked_indices = ranked_indices.cpu().float().numpy()                           # :362
    pred_dicts = []
    for ranked_idx, gt_info in zip(ranked_indices, gt_infos):                           # :367
        rel_docids = gt_info["label_name"] if isinstance(gt_info["label_name"], list) else [gt_info["label_name"]]
        pred_dicts.append({"prediction": [cand_keys[int(i)] for i in ranked_idx], "label": rel_docids,
                           "rel_scores": gt_info.get("rel_scores", None)})
    return scores.numpy(), ranked_indices.astype(np.int64), pred_dicts


def run_local(qry_embeds, cand_embed_dict, gt_infos):
    qry_tensor = torch.from_numpy(qry_embeds)
    pred_dicts, all_scores, all_ranked = [], [], []
    for qry_vec, gt_info in zip(qry_tensor, gt_infos):                                  # :384
        cand_names = gt_info["cand_names"]                                              # :389
        cand_embeds = np.stack([cand_embed_dict[name] for name in cand_names])          # :390
        cand_tensor = torch.from_numpy(cand_embeds)                                     # :391
        with torch.no_grad():
            sim_scores = compute_similarity(qry_vec.unsqueeze(0), cand_tensor).squeeze(0)   # :394
            _, ranked_idx = torch.sort(sim_scores, descending=True)                     # :395
            ranked_idx = ranked_idx.cpu().float().numpy()                               # :396
        rel_docids = gt_info["label_name"] if isinstance(gt_info["label_name"], list) else [gt_info["label_name"]]
        pred_dicts.append({"prediction": [cand_names[int(i)] for i in ranked_idx], "label": rel_docids,
                           "rel_scores": gt_info.get("rel_scores", None)})
        all_scores.append(sim_scores.numpy())
        all_ranked.append(ranked_idx.astype(np.int64))
    return all_scores, all_ranked, pred_dicts


def truncated(pred_dicts, k):
    return [dict(p, prediction=p["prediction"][:k]) for p in pred_dicts]


def main():
    out = {}
    for name in MMEB_CASES:
        qry, cdict, gts = gen_mmeb_case(name)
        out[f"{name}/qry_sha"] = np.frombuffer(bytes.fromhex(sha(qry)), dtype=np.uint8)
        out[f"{name}/cand_sha"] = np.frombuffer(bytes.fromhex(sha(np.stack(list(cdict.values())))), dtype=np.uint8)
        scores, ranked, preds = run_global(qry, cdict, gts)
        m_full = RankingMetrics(METRICS).evaluate(preds)
        m_top10 = RankingMetrics(METRICS).evaluate(truncated(preds, 10))
        assert m_full == m_top10, "metrics must only read prediction[:10]"
        out[f"{name}/global/scores"] = scores
        out[f"{name}/global/ranked"] = ranked
        out[f"{name}/global/metrics"] = np.frombuffer(json.dumps(m_full, sort_keys=True).encode(), dtype=np.uint8)
        lscores, lranked, lpreds = run_local(qry, cdict, gts)
        lm_full = RankingMetrics(METRICS).evaluate(lpreds)
        assert lm_full == RankingMetrics(METRICS).evaluate(truncated(lpreds, 10))
        out[f"{name}/local/offsets"] = np.concatenate([[0], np.cumsum([len(r) for r in lranked])]).astype(np.int64)
        out[f"{name}/local/scores"] = np.concatenate(lscores)
        out[f"{name}/local/ranked"] = np.concatenate(lranked)
        out[f"{name}/local/metrics"] = np.frombuffer(json.dumps(lm_full, sort_keys=True).encode(), dtype=np.uint8)
        print(name, "global hit@1", m_full["hit@1"], "local hit@1", lm_full["hit@1"])
    path = os.path.join(HERE, "mmeb_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
