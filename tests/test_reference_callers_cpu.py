"""CPU half of "run the reference's callers on the engine" (rows a7/a8/a9): with /root/reference importable in this container,
``nexus_b200.patch.install()`` is applied and the reference's REAL callers are run end to end --

  * ``TextRetrievalEvalDenseRetriever.__call__``        Nexus/evaluation/text_retrieval/searcher.py:77-161
  * ``MultimodalRetrievalEvalDenseRetriever.__call__``  Nexus/evaluation/multimodal_retrieval/searcher.py:35-73
  * ``BaseRetriever.eval_step``                         Nexus/training/embedder/recommendation/modeling.py:176-207
  * ``index()`` with the faiss branch taken             Nexus/evaluation/text_retrieval/utils.py:174-191

-- with a stub embedder and the ctypes layer replaced by an oracle-backed fake (tests/_fake_lib.py; there is no GPU here), and
compared with the same callers UNPATCHED (the reference's numpy fallback).  What this pins: the patched names are the ones the
callers hit, the arguments they pass are accepted (``index=`` by keyword, ``GpuMultipleClonerOptions``), and the result objects
have the reference's shape and contents.  The GPU half (tests/test_gpu_callers.py) replays the same statements on the real library.
The tests skip where the reference tree is absent (the GPU box)."""
import ast
import importlib
import importlib.machinery
import os
import sys
import types

import numpy as np
import pytest

REF = "/root/reference"
pytestmark = pytest.mark.skipif(not os.path.isdir(os.path.join(REF, "Nexus")), reason="reference tree not present")


class _Stub(types.ModuleType):
    def __init__(self, name):
        super().__init__(name)
        self.__spec__ = importlib.machinery.ModuleSpec(name, None)
        self.__path__ = []

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        m = _Stub(self.__name__ + "." + name)
        setattr(self, name, m)
        return m

    def __call__(self, *a, **k):
        return self


@pytest.fixture
def engine(monkeypatch):
    """Reference importable (the four absent optional imports stubbed, as tests/golden/make_golden.py does), a pretend CUDA
    device, the fake library behind nexus_b200._lib.load(); everything is undone afterwards."""
    import torch
    from _fake_lib import FakeLib
    from nexus_b200 import _lib, patch
    added_path = REF not in sys.path
    if added_path:
        sys.path.insert(0, REF)
    stubbed = [n for n in ["onnx", "onnxruntime", "tensorrt", "pycuda", "pycuda.driver", "pycuda.autoinit"] if n not in sys.modules]
    for n in stubbed:
        sys.modules[n] = _Stub(n)
    saved_faiss = {k: sys.modules.get(k) for k in ("faiss", "faiss.contrib", "faiss.contrib.torch_utils")}
    fake = FakeLib()
    monkeypatch.setattr(_lib, "load", lambda: fake)
    monkeypatch.setattr(torch.cuda, "is_available", lambda: True)
    monkeypatch.setattr(torch.cuda, "device_count", lambda: 1)
    monkeypatch.setattr(torch.cuda, "current_device", lambda: 0)
    monkeypatch.setattr(torch.cuda, "empty_cache", lambda: None)
    yield fake, patch
    patch.uninstall()
    for k, v in saved_faiss.items():
        if v is None:
            sys.modules.pop(k, None)
        else:
            sys.modules[k] = v
    for n in stubbed:
        sys.modules.pop(n, None)
    if added_path:
        sys.path.remove(REF)


class StubEmbedder:
    """encode_info / encode_query (text) and encode_corpus / encode_queries (multimodal): deterministic unit vectors per item."""

    def __init__(self, dim=48):
        self.dim = dim
        self.calls = []

    def _emb(self, items, salt):
        out = np.empty((len(items), self.dim), np.float32)
        for i, it in enumerate(items):
            rng = np.random.default_rng(abs(hash((salt, str(it)))) % (2 ** 32))
            v = rng.standard_normal(self.dim).astype(np.float32)
            out[i] = v / np.linalg.norm(v)
        return out

    def encode_info(self, texts, **kw):
        self.calls.append(("info", len(texts)))
        return self._emb(texts, 0)

    def encode_query(self, texts, **kw):
        self.calls.append(("query", len(texts)))
        return self._emb(texts, 0)          # same map as the corpus: a query equal to a document text retrieves it first

    encode_corpus = encode_info
    encode_queries = encode_query


def _text_case():
    corpus = {f"doc-{i}": ({"text": f"passage number {i}"} if i % 3 else {"title": f"T{i}", "text": f"passage number {i}"}) for i in range(300)}
    queries = {f"q-{i}": f"passage number {i * 7 % 300}" for i in range(40)}
    queries["doc-5"] = "passage number 5"        # a query id that IS a corpus id (ignore_identical_ids)
    return corpus, queries


def _assert_same_results(got, ref):
    assert list(got.keys()) == list(ref.keys())
    for qid in ref:
        assert list(got[qid].keys()) == list(ref[qid].keys()), qid          # same documents in the same (insertion) order
        for d, v in ref[qid].items():
            assert type(got[qid][d]) is float and got[qid][d] == v


@pytest.mark.parametrize("ignore", [False, True])
@pytest.mark.parametrize("top_k", [10, 1000])           # 1000 > 300 documents: the engine pads with -1, the numpy fallback truncates
def test_text_dense_retriever_call_patched_equals_unpatched(engine, tmp_path, ignore, top_k):
    fake, patch = engine
    searcher = importlib.import_module("Nexus.evaluation.text_retrieval.searcher")
    utils = importlib.import_module("Nexus.evaluation.text_retrieval.utils")
    corpus, queries = _text_case()
    r = searcher.TextRetrievalEvalDenseRetriever.__new__(searcher.TextRetrievalEvalDenseRetriever)
    r.embedder, r.search_top_k, r.overwrite = StubEmbedder(), top_k, False
    assert utils.faiss is None                                  # unpatched: the reference's numpy fallback (utils.py:170-172)
    ref = r(corpus, queries, corpus_embd_save_dir=str(tmp_path / "ref"), ignore_identical_ids=ignore)
    done = patch.install()
    assert done["Nexus.evaluation.text_retrieval.searcher"] == "patched"
    assert done["Nexus.evaluation.text_retrieval.searcher.TextRetrievalEvalDenseRetriever"] == "patched"
    import nexus_b200.retrieval as R
    assert searcher.index is R.index and searcher.search is R.search and utils.index is R.index
    got = r(corpus, queries, corpus_embd_save_dir=str(tmp_path / "got"), ignore_identical_ids=ignore)
    assert fake.calls["create"] == 1 and fake.calls["add"] == 1 and fake.calls["search"] == 1      # ONE search call, not 32-row chunks
    assert fake.calls["search_rows"] == len(queries)
    _assert_same_results(got, ref)
    assert np.array_equal(np.load(tmp_path / "got" / "doc.npy"), np.load(tmp_path / "ref" / "doc.npy"))      # searcher.py:141-144
    # second call: doc.npy is read back instead of encoding the corpus again (searcher.py:126-127)
    n_info = sum(1 for c in r.embedder.calls if c[0] == "info")
    got2 = r(corpus, queries, corpus_embd_save_dir=str(tmp_path / "got"), ignore_identical_ids=ignore)
    assert sum(1 for c in r.embedder.calls if c[0] == "info") == n_info
    _assert_same_results(got2, ref)
    # the reference's own __call__ body on the patched index/search (retriever_call=False): its loop skips the -1 padding
    patch.uninstall()
    patch.install(retriever_call=False)
    got3 = r(corpus, queries, ignore_identical_ids=ignore)
    _assert_same_results(got3, ref)
    patch.uninstall()
    assert searcher.index is not R.index


def test_multimodal_dense_retriever_call_patched_equals_unpatched(engine, tmp_path):
    fake, patch = engine
    mm = importlib.import_module("Nexus.evaluation.multimodal_retrieval.searcher")
    corpus = {f"img-{i}": {"text": f"caption {i}"} for i in range(120)}
    queries = {f"q-{i}": {"text": f"caption {i * 5 % 120}"} for i in range(25)}
    queries["img-3"] = {"text": "caption 3"}
    r = mm.MultimodalRetrievalEvalDenseRetriever.__new__(mm.MultimodalRetrievalEvalDenseRetriever)
    r.embedder, r.search_top_k, r.overwrite = StubEmbedder(32), 20, False
    ref = r(corpus, queries, ignore_identical_ids=True)
    patch.install(retriever_call=False)             # the method imports index/search from utils at call time (searcher.py:43)
    got_loop = r(corpus, queries, ignore_identical_ids=True)
    assert fake.calls["search"] == 1
    _assert_same_results(got_loop, ref)
    patch.uninstall()
    patch.install()
    got = r(corpus, queries, corpus_embd_save_dir=str(tmp_path / "mm"), ignore_identical_ids=True)
    _assert_same_results(got, ref)
    assert os.path.exists(tmp_path / "mm" / "corpus_embeddings.npy")


def test_index_takes_the_faiss_gpu_branch_on_the_shim(engine):
    """The reference's OWN index()/search() (not our replacements) with ``faiss`` = the shim: index_factory -> GpuMultipleClonerOptions
    (shard, useFloat16) -> index_cpu_to_all_gpus -> train -> add -> 32-row search chunks (utils.py:174-229)."""
    fake, patch = engine
    from nexus_b200 import faiss_shim
    utils = importlib.import_module("Nexus.evaluation.text_retrieval.utils")
    ref_index, ref_search = utils.index, utils.search
    c = StubEmbedder(64)._emb([f"d{i}" for i in range(500)], 1)
    q = StubEmbedder(64)._emb([f"d{i}" for i in range(0, 500, 7)], 1)
    base_s, base_i = ref_search(ref_index(corpus_embeddings=c), k=10, query_embeddings=q)           # numpy fallback
    patch.install()
    assert utils.faiss is sys.modules["faiss"] and getattr(utils.faiss, "__nexus_b200__", False)
    saved = utils._faiss_supports_gpu if hasattr(utils, "_faiss_supports_gpu") else None
    try:
        if saved is not None:
            utils._faiss_supports_gpu = lambda: True
        idx = ref_index(corpus_embeddings=c)                    # the reference's function object, faiss branch
        assert isinstance(idx, faiss_shim._LazyFlat) and idx._store_dtype == "f16" and idx._devices == [0]      # co.useFloat16, all GPUs
        s, i = ref_search(idx, k=10, query_embeddings=q)        # the reference's 32-row loop
    finally:
        if saved is not None:
            utils._faiss_supports_gpu = saved
    assert fake.calls["search"] == -(-q.shape[0] // 32)
    assert s.shape == base_s.shape and i.dtype == np.int64
    np.testing.assert_array_equal(i[:, 0], base_i[:, 0])        # fp16 storage: the self-match stays first
    assert np.abs(s - base_s).max() < 2e-3


def _reference_eval_step():
    """``BaseRetriever.eval_step`` compiled from the reference's source text, unmodified (importing its module pulls accelerate /
    torchrec, absent here); its globals are torch, Tuple and ``faiss`` = whatever sys.modules holds."""
    import torch
    from typing import Tuple
    path = os.path.join(REF, "Nexus/training/embedder/recommendation/modeling.py")
    tree = ast.parse(open(path).read())
    cls = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == "BaseRetriever")
    fn = next(n for n in cls.body if isinstance(n, ast.FunctionDef) and n.name == "eval_step")
    mod = ast.Module(body=[fn], type_ignores=[])
    ns = {"torch": torch, "Tuple": Tuple, "faiss": sys.modules["faiss"]}
    exec(compile(mod, path, "exec"), ns)
    return ns["eval_step"]


class _Dev:
    class device:
        index = 0


class InnerProductScorer:
    def __call__(self, q, i):
        return (q * i).sum(-1)


class EuclideanScorer:
    def __call__(self, q, i):
        return -((q - i) ** 2).sum(-1)         # modules/score.py: larger is better; Faiss L2 returns ascending squared distances


class CosineScorer:
    def __call__(self, q, i):
        import torch
        return (torch.nn.functional.normalize(q, dim=-1) * torch.nn.functional.normalize(i, dim=-1)).sum(-1)


class _Retriever:
    fiid = "item_id"

    def __init__(self, scorer, qv, pv):
        self.score_function, self._qv, self._pv = scorer, qv, pv
        outer = self

        class Enc:
            def __call__(self, batch):
                return outer._qv

            def parameters(self):
                return iter([_Dev()])
        self.query_encoder = Enc()
        self.item_encoder = lambda batch: outer._pv


@pytest.mark.parametrize("scorer", [InnerProductScorer, CosineScorer])
def test_eval_step_runs_on_the_shim(engine, scorer):
    """modeling.py:176-207 verbatim on the shim: IndexFlatIP(d) -> index_cpu_to_gpu(res, gpu_index, index=index_flat) BY KEYWORD ->
    add(item_vectors) -> search(query_vec, k) with torch tensors in and out -> cat + stable sort -> pred."""
    import torch
    import oracle
    fake, patch = engine
    patch.install()
    eval_step = _reference_eval_step()
    gen = torch.Generator().manual_seed(3)
    items = torch.randn(400, 16, generator=gen)
    target = torch.randint(0, 400, (32,), generator=gen)
    qv = torch.randn(32, 16, generator=gen) + 1.5 * items[target]
    pos_vec = items[target] + 0.05 * torch.randn(32, 16, generator=gen)     # near its item, never bit-equal to a table row: no rounding ties
    model = _Retriever(scorer(), qv, pos_vec)
    batch = {"item_id": target}
    pred, tgt = eval_step(model, batch, 10, items)
    assert pred.dtype == torch.bool and tuple(pred.shape) == (32, 10) and tuple(tgt.shape) == (32, 1) and bool(tgt.all())
    it = torch.nn.functional.normalize(items, dim=-1).numpy() if scorer is CosineScorer else items.numpy()
    pv = torch.nn.functional.normalize(pos_vec, dim=-1).numpy() if scorer is CosineScorer else pos_vec.numpy()
    pos = np.asarray(scorer()(qv, torch.from_numpy(pv)))
    top_s, _ = oracle.blocked_ip_search(it, qv.numpy(), 10)
    rank, ref_pred = oracle.positive_pred(pos, top_s)
    np.testing.assert_array_equal(pred.numpy(), ref_pred)
    if scorer is InnerProductScorer:     # (Cosine: the reference scores the positive with normalised queries and the table with raw
        assert (rank < 10).sum() > 0     #  ones, modeling.py:181,188-189,197 -- the positives rank last; reproduced, not judged)
    assert fake.calls["create"] == 1 and fake.calls["add"] == 1 and fake.calls["search"] == 1
