"""Drop-in ``index()`` / ``search()`` with the reference's signatures and return shapes.

Mirrors Nexus/evaluation/text_retrieval/utils.py:148-191 (``index``) and :194-230 (``search``), the two
functions the dense retrievers call (text_retrieval/searcher.py:149-150, multimodal_retrieval/searcher.py:62-63).
Differences, all deliberate:
  * the index lives in HBM on a B200 and every search runs in hand-written sm_100a kernels; there is no
    CPU Faiss / numpy fallback (the reference silently downgrades at utils.py:170-172,184-185);
  * ``search`` hands the WHOLE query matrix to the engine in one call instead of 32 rows at a time
    (utils.py:221-226), which is what lets the tensor-core path run;
  * results are Faiss-shaped ``[Q, k]`` with ``-1`` padding when k > N (the numpy fallback truncates
    columns, utils.py:23; callers already skip ``-1``, searcher.py:156).
"""
from __future__ import annotations

import logging
import os
from typing import Any, Optional, Tuple

import numpy as np

from .flat import FlatIndex, parse_device
from .sharded import ShardedFlatIndex

logger = logging.getLogger(__name__)


def _default_store_dtype(emb_dtype) -> str:
    """Storage type when the caller does not choose: keep what the embedder produced when that is
    already 16-bit (``use_fp16`` encoders hand fp16 arrays, base.py:272-273), keep fp32 exact
    otherwise so that scores match the reference's fp32 arithmetic to 1e-5.  Override with
    ``NEXUS_B200_STORE_DTYPE=bf16|f16|f32`` (bf16/f16 select the tcgen05 path for fp32 inputs too)."""
    env = os.environ.get("NEXUS_B200_STORE_DTYPE")
    if env:
        return env
    s = str(emb_dtype)
    if "bfloat16" in s:
        return "bf16"
    if "float16" in s:
        return "f16"
    return "f32"


def index(index_factory: str = "Flat", corpus_embeddings: Optional[np.ndarray] = None, load_path: Optional[str] = None,
          device: Optional[str] = None, store_dtype: Optional[str] = None, shard: Optional[bool] = None):
    """Build a flat inner-product index and add the corpus embeddings (utils.py:148-191).

    ``device=None`` follows the reference (utils.py:178-183): use every visible GPU, rows sharded
    (``co.shard = True``); with one visible GPU that is a single :class:`FlatIndex`.  ``device='cuda:1'``
    pins one device.  ``corpus_embeddings`` may be a numpy array or a torch tensor (CPU or CUDA -- a CUDA
    tensor is ingested without leaving the device)."""
    if corpus_embeddings is None:
        corpus_embeddings = np.load(load_path)             # utils.py:165-166
    if index_factory != "Flat":
        raise ValueError(f"index_factory={index_factory!r}: this engine builds exact flat indices only ('Flat')")
    logger.info(f"Shape of embeddings: {tuple(corpus_embeddings.shape)}")
    d = int(corpus_embeddings.shape[-1])
    sd = store_dtype or _default_store_dtype(corpus_embeddings.dtype)
    import torch
    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n_gpu == 0:
        raise RuntimeError("nexus_b200.index: no CUDA device is visible and this engine has no CPU fallback")
    on_dev = getattr(corpus_embeddings, "is_cuda", False)
    if device is None and not on_dev and (shard if shard is not None else n_gpu > 1):
        idx = ShardedFlatIndex(d, "ip", sd, devices=list(range(n_gpu)))
    else:
        dev = corpus_embeddings.device.index if on_dev and device is None else parse_device(device)
        idx = FlatIndex(d, "ip", sd, device=dev)
    logger.info(f"Indexing {corpus_embeddings.shape[0]} documents...")
    idx.train(corpus_embeddings)                            # utils.py:188 (no-op for Flat)
    idx.add(corpus_embeddings)                              # utils.py:189
    logger.info("Embeddings add over...")
    return idx


def search(faiss_index: Any, k: int = 100, query_embeddings: Optional[np.ndarray] = None, load_path: Optional[str] = None
           ) -> Tuple[np.ndarray, np.ndarray]:
    """Search the index (utils.py:194-230): returns ``(all_scores float32 [Q,k], all_indices int64 [Q,k])``."""
    if query_embeddings is None:
        query_embeddings = np.load(load_path)               # utils.py:212-213
    # every engine-backed object takes the WHOLE query matrix in one call: FlatIndex / ShardedFlatIndex, the shim's lazily
    # materialised Faiss-shaped objects (index_factory, read_index, index_cpu_to_all_gpus) and a DistributedFlatIndex
    engine = faiss_index
    if getattr(faiss_index, "__nexus_b200_lazy__", False):
        engine = faiss_index._ensure(query_embeddings)
    if isinstance(engine, (FlatIndex, ShardedFlatIndex)) or getattr(engine, "__nexus_b200_engine__", False):
        scores, ids = engine.search(query_embeddings, k)
        if not isinstance(scores, np.ndarray):              # torch in -> torch out; this function returns numpy
            scores, ids = scores.cpu().numpy(), ids.cpu().numpy()
        return scores, ids
    # a foreign index object (e.g. a real faiss index): keep the reference's host loop (utils.py:221-229)
    all_scores, all_indices = [], []
    for i in range(0, len(query_embeddings), 32):
        s, ix = faiss_index.search(np.asarray(query_embeddings[i:i + 32]).astype(np.float32), k)
        all_scores.append(s)
        all_indices.append(ix)
    return np.concatenate(all_scores, axis=0), np.concatenate(all_indices, axis=0)


def unpack_results(queries_ids, corpus_ids, all_scores, all_indices, ignore_identical_ids: bool = False):
    """``{query_id: {doc_id: float(score)}}`` from the ``[Q, k]`` arrays, identical to the nested loop at
    Nexus/evaluation/text_retrieval/searcher.py:152-159 (``-1`` ids skipped; with ``ignore_identical_ids`` a document whose
    id equals the query's is dropped) but without Q*k Python-level iterations: the arrays are converted once
    (``tolist``), ids are looked up per row with list indexing, and rows without padding take a dict(zip()) fast path.
    Later duplicates of a document id overwrite earlier ones, as in the reference."""
    scores = np.asarray(all_scores)
    indices = np.asarray(all_indices)
    if scores.shape != indices.shape or scores.ndim != 2:
        raise ValueError(f"scores {scores.shape} and indices {indices.shape} must be the same [Q, k] shape")
    if len(queries_ids) != scores.shape[0]:
        raise ValueError(f"{len(queries_ids)} query ids for {scores.shape[0]} result rows")
    corpus_ids = corpus_ids if isinstance(corpus_ids, list) else list(corpus_ids)
    has_pad = (indices < 0).any(axis=1).tolist()
    s_rows = scores.astype(np.float64).tolist()       # float(np.float32) == the float64 value of the float32
    i_rows = indices.tolist()
    results = {}
    for qid, srow, irow, pad in zip(queries_ids, s_rows, i_rows, has_pad):
        if pad:
            pairs = [(corpus_ids[i], s) for i, s in zip(irow, srow) if i != -1]
        else:
            pairs = list(zip([corpus_ids[i] for i in irow], srow))
        if ignore_identical_ids:
            pairs = [(d, s) for d, s in pairs if d != qid]
        results[qid] = dict(pairs)
    return results
