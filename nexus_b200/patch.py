"""Switch a running Nexus process over to the B200 engine.

``install()`` rebinds the names the dense retrievers call:
  * ``Nexus.evaluation.text_retrieval.utils.index`` / ``.search``   (looked up at call time by
    multimodal_retrieval/searcher.py:43,62-63)
  * ``Nexus.evaluation.text_retrieval.searcher.index`` / ``.search`` (bound at import time, searcher.py:14)
  * ``TextRetrievalEvalDenseRetriever.__call__`` (text_retrieval/searcher.py:77-161) and
    ``MultimodalRetrievalEvalDenseRetriever.__call__`` (multimodal_retrieval/searcher.py:35-73): same arguments, same cache
    files, same result dictionary, but the ``Q*k`` Python-level loop that builds it (searcher.py:152-159 / :65-73) is replaced
    by :func:`nexus_b200.retrieval.unpack_results` (row a8; ``retriever_call=False`` leaves the methods alone)
and, when no real Faiss is importable, registers :mod:`nexus_b200.faiss_shim` as ``faiss`` so that
``BaseRetriever.eval_step`` (training/embedder/recommendation/modeling.py:176-207), the recsys engines
(inference/embedder/recommendation/base.py:114-117,182) and the serve script
(serve/inference_service_text_retrieval.py:35,70) run on it unchanged.
"""
from __future__ import annotations

import gc
import importlib
import os
import sys

import numpy as np

from . import faiss_shim, retrieval

_saved = {}


def text_dense_retriever_call(self, corpus, queries, corpus_embd_save_dir=None, ignore_identical_ids=False, **kwargs):
    """``TextRetrievalEvalDenseRetriever.__call__`` (text_retrieval/searcher.py:77-161) on the engine: the embedder calls, the
    ``doc.npy`` cache and the returned ``{qid: {docid: float(score)}}`` are the reference's; index / search run on the B200 and the
    dictionary is built by ``unpack_results`` instead of ``Q*k`` Python iterations."""
    import torch
    kwargs.pop("language", None)                                              # searcher.py:108
    corpus_ids = list(corpus.keys())
    corpus_texts = [doc["text"] if "title" not in doc else f"{doc['title']} {doc['text']}".strip() for doc in corpus.values()]   # :112-117
    queries_ids = list(queries.keys())
    queries_texts = list(queries.values())
    doc_path = os.path.join(corpus_embd_save_dir, "doc.npy") if corpus_embd_save_dir is not None else None
    if doc_path is not None and os.path.exists(doc_path) and not self.overwrite:
        corpus_emb = np.load(doc_path)                                        # :126-127
        saved = True
    else:
        corpus_emb = self.embedder.encode_info(corpus_texts, **kwargs)        # :129,131
        saved = False
    queries_emb = self.embedder.encode_query(queries_texts, **kwargs)         # :133
    if isinstance(corpus_emb, dict):
        corpus_emb = corpus_emb["dense_vecs"]                                 # :136-139 (M3Embedder)
    if isinstance(queries_emb, dict):
        queries_emb = queries_emb["dense_vecs"]
    if doc_path is not None and not saved:
        os.makedirs(corpus_embd_save_dir, exist_ok=True)
        np.save(doc_path, corpus_emb)                                         # :141-144
    gc.collect()
    if torch.cuda.is_available():
        torch.cuda.empty_cache()                                              # :146-147
    faiss_index = retrieval.index(corpus_embeddings=corpus_emb)               # :149
    all_scores, all_indices = retrieval.search(query_embeddings=queries_emb, faiss_index=faiss_index, k=self.search_top_k)   # :150
    return retrieval.unpack_results(queries_ids, corpus_ids, all_scores, all_indices, ignore_identical_ids)                   # :152-159


def multimodal_dense_retriever_call(self, corpus, queries, corpus_embd_save_dir=None, ignore_identical_ids=False, **kwargs):
    """``MultimodalRetrievalEvalDenseRetriever.__call__`` (multimodal_retrieval/searcher.py:35-73) on the engine, same contract."""
    corpus_ids = list(corpus.keys())
    query_ids = list(queries.keys())
    path = None
    if corpus_embd_save_dir is not None:
        os.makedirs(corpus_embd_save_dir, exist_ok=True)
        path = os.path.join(corpus_embd_save_dir, "corpus_embeddings.npy")    # :48-52
    if path is not None and os.path.exists(path) and not self.overwrite:
        corpus_embeddings = np.load(path)                                     # :54-55
    else:
        corpus_embeddings = self.embedder.encode_corpus([corpus[cid] for cid in corpus_ids])
        if path is not None:
            np.save(path, corpus_embeddings)                                  # :57-59
    query_embeddings = self.embedder.encode_queries([queries[qid] for qid in query_ids])
    faiss_index = retrieval.index(corpus_embeddings=corpus_embeddings)        # :62
    all_scores, all_indices = retrieval.search(faiss_index=faiss_index, k=self.search_top_k, query_embeddings=query_embeddings)
    return retrieval.unpack_results(query_ids, corpus_ids, all_scores, all_indices, ignore_identical_ids)   # :65-73


_RETRIEVER_CALLS = (
    ("Nexus.evaluation.text_retrieval.searcher", "TextRetrievalEvalDenseRetriever", text_dense_retriever_call),
    ("Nexus.evaluation.multimodal_retrieval.searcher", "MultimodalRetrievalEvalDenseRetriever", multimodal_dense_retriever_call),
)


def install(faiss: bool = True, retriever_call: bool = True) -> dict:
    done = {}
    if faiss:
        mod = faiss_shim.install_as_faiss()
        done["faiss"] = getattr(mod, "__nexus_b200__", False)
    for modname in ("Nexus.evaluation.text_retrieval.utils", "Nexus.evaluation.text_retrieval.searcher"):
        try:
            m = sys.modules.get(modname) or importlib.import_module(modname)
        except Exception as e:  # heavy optional deps of the reference (onnx, tensorrt, ...) may be absent
            done[modname] = f"not patched: {type(e).__name__}"
            continue
        for name in ("index", "search"):
            if hasattr(m, name):
                _saved.setdefault((modname, name), getattr(m, name))
                setattr(m, name, getattr(retrieval, name))
        if hasattr(m, "faiss") and faiss and done.get("faiss"):
            _saved.setdefault((modname, "faiss"), m.faiss)
            m.faiss = sys.modules["faiss"]
        done[modname] = "patched"
    if retriever_call:
        for modname, clsname, fn in _RETRIEVER_CALLS:
            try:
                m = sys.modules.get(modname) or importlib.import_module(modname)
                cls = getattr(m, clsname)
            except Exception as e:  # noqa: BLE001
                done[f"{modname}.{clsname}"] = f"not patched: {type(e).__name__}"
                continue
            _saved.setdefault((modname, clsname), cls.__call__)
            cls.__call__ = fn
            done[f"{modname}.{clsname}"] = "patched"
    return done


def uninstall() -> None:
    for (modname, name), fn in list(_saved.items()):
        m = sys.modules.get(modname)
        if m is not None:
            if isinstance(getattr(m, name, None), type):     # a class whose __call__ was replaced
                getattr(m, name).__call__ = fn
            else:
                setattr(m, name, fn)
        _saved.pop((modname, name))
