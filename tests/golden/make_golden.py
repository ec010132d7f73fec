"""Generate golden vectors by running the REFERENCE's own code in the build container.

Run once here (``python tests/golden/make_golden.py``); ``/root/reference`` does not exist on the
GPU box, so the outputs are committed as small ``.npz`` fixtures next to this script and nothing
in the test-suite reads ``/root/reference`` at run time.

What is executed, unmodified, from ``/root/reference``:
  * ``Nexus.evaluation.text_retrieval.utils.index`` / ``search`` / ``NumpyIPIndex`` (utils.py:16-31,148-230)
  * ``BaseEmbedder.pooling`` (inference/embedder/text_retrieval/base.py:355-379)
  * ``BaseDecoderEmbedder.pooling`` (inference/embedder/text_retrieval/decoder/base.py:330-342)
  * ``MultimodalEmbedder._pool_hidden_states`` (inference/embedder/multimodal_retrieval/generic.py:104-126)
  * ``MMEBEmbeddingModel._pooling_last`` (evaluation/mmeb_v2/models.py:146-151)
  * ``torch.nn.functional.normalize`` as called at base.py:334-335
The four heavy optional imports those modules make at import time (onnx, onnxruntime, tensorrt,
pycuda) are absent from this image and are stubbed in ``sys.modules``; none is on the path exercised.
"""
from __future__ import annotations

import hashlib
import importlib
import importlib.machinery
import os
import sys
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"


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


def _load_reference():
    sys.path.insert(0, REF)
    for name in ["onnx", "onnxruntime", "tensorrt", "pycuda", "pycuda.driver", "pycuda.autoinit"]:
        sys.modules.setdefault(name, _Stub(name))
    utils = importlib.import_module("Nexus.evaluation.text_retrieval.utils")
    assert utils.faiss is None, "goldens are defined by the numpy fallback; faiss must be absent"
    enc = importlib.import_module("Nexus.inference.embedder.text_retrieval.base")
    dec = importlib.import_module("Nexus.inference.embedder.text_retrieval.decoder.base")
    mm = importlib.import_module("Nexus.inference.embedder.multimodal_retrieval.generic")
    mmeb = importlib.import_module("Nexus.evaluation.mmeb_v2.models")
    return utils, enc, dec, mm, mmeb


def sha(a: np.ndarray) -> str:
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gen_inputs(n, d, q, seed, normalize=True, dtype=np.float32):
    """The input recipe shared with tests/: numpy PCG64, corpus from ``seed``, queries from ``seed+1000``."""
    rc = np.random.default_rng(seed)
    rq = np.random.default_rng(seed + 1000)
    c = rc.standard_normal((n, d), dtype=np.float32)
    x = rq.standard_normal((q, d), dtype=np.float32)
    if normalize:
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    return c.astype(dtype), x.astype(dtype)


SEARCH_CASES = [
    # name, N, D, Q, k, seed, normalize, dtype, store_inputs
    ("tiny_ragged", 257, 32, 37, 10, 11, True, np.float32, True),
    ("bge_small_slice", 1000, 384, 70, 10, 12, True, np.float32, False),
    ("k_gt_n", 7, 16, 5, 10, 13, True, np.float32, True),
    ("k_eq_n", 16, 16, 9, 16, 14, True, np.float32, True),
    ("k_one", 300, 64, 33, 1, 15, True, np.float32, True),
    ("fp16_inputs", 500, 128, 40, 20, 16, True, np.float16, False),
    ("unnormalised_d64", 2000, 64, 64, 50, 17, False, np.float32, False),
    ("k100", 5000, 96, 45, 100, 18, True, np.float32, False),
    ("k1000", 4096, 48, 17, 1000, 19, True, np.float32, False),
    ("odd_dim", 777, 100, 21, 7, 20, True, np.float32, False),
    ("c1_bge_small", 100000, 384, 1000, 10, 1001, True, np.float32, False),  # BASELINE config 1
]


def main():
    import torch

    utils, enc, dec, mm, mmeb = _load_reference()
    out = {}

    # ---------------- search: reference index() + search() (numpy fallback) ----------------
    for name, n, d, q, k, seed, norm, dt, store in SEARCH_CASES:
        c, x = gen_inputs(n, d, q, seed, norm, dt)
        idx = utils.index(corpus_embeddings=c)
        s, i = utils.search(idx, k=k, query_embeddings=x)
        assert s.dtype == np.float32 and i.dtype == np.int64
        out[f"search/{name}/meta"] = np.array([n, d, q, k, seed, int(norm), 1 if dt == np.float16 else 0], dtype=np.int64)
        out[f"search/{name}/scores"] = s
        out[f"search/{name}/ids"] = i
        out[f"search/{name}/corpus_sha"] = np.frombuffer(bytes.fromhex(sha(c)), dtype=np.uint8)
        out[f"search/{name}/query_sha"] = np.frombuffer(bytes.fromhex(sha(x)), dtype=np.uint8)
        if store:
            out[f"search/{name}/corpus"] = c
            out[f"search/{name}/queries"] = x
        print(f"search/{name}: N={n} D={d} Q={q} k={k} -> {s.shape}")

    # known-answer: identity corpus (KAT-1) and duplicate rows (KAT-2), through the reference
    d = 24
    c = np.eye(d, dtype=np.float32)
    x = (np.eye(d, dtype=np.float32)[[3, 7, 23]] * 2.0 + 0.01).astype(np.float32)
    s, i = utils.search(utils.index(corpus_embeddings=c), k=3, query_embeddings=x)
    out["kat/identity/corpus"], out["kat/identity/queries"] = c, x
    out["kat/identity/scores"], out["kat/identity/ids"] = s, i
    rng = np.random.default_rng(5)
    base = rng.standard_normal((8, 16), dtype=np.float32)
    c = np.concatenate([base, base, base[:3]], axis=0)          # rows j, j+8 (and j+16 for j<3) identical
    x = rng.standard_normal((6, 16), dtype=np.float32)
    s, i = utils.search(utils.index(corpus_embeddings=c), k=6, query_embeddings=x)
    out["kat/duplicates/corpus"], out["kat/duplicates/queries"] = c, x
    out["kat/duplicates/scores"], out["kat/duplicates/ids"] = s, i

    # the reference's only fixture shape: 2 documents, 1 query
    # (examples/multimodal_retrieval/data/eval/{corpus,test_queries}.jsonl)
    c, x = gen_inputs(2, 8, 1, 21)
    s, i = utils.search(utils.index(corpus_embeddings=c), k=10, query_embeddings=x)
    out["kat/toy/corpus"], out["kat/toy/queries"], out["kat/toy/scores"], out["kat/toy/ids"] = c, x, s, i

    # ---------------- pooling + normalise ----------------
    g = torch.Generator().manual_seed(77)
    B, L, H = 6, 9, 40
    hidden = torch.randn(B, L, H, generator=g)
    lens = [9, 1, 4, 7, 2, 5]
    right = torch.zeros(B, L, dtype=torch.int64)
    left = torch.zeros(B, L, dtype=torch.int64)
    for b, n_ in enumerate(lens):
        right[b, :n_] = 1
        left[b, L - n_:] = 1
    holes = right.clone()
    holes[0, 3] = 0
    holes[3, 1] = 0
    holes[3, 2] = 0
    full = torch.ones(B, L, dtype=torch.int64)
    masks = {"right": right, "left": left, "holes": holes, "full": full}
    out["pool/hidden"] = hidden.numpy()
    for mname, m in masks.items():
        out[f"pool/mask_{mname}"] = m.numpy()

    def norm(t):
        return torch.nn.functional.normalize(t, dim=-1)

    for mname, m in masks.items():
        for pm in ["cls", "mean"]:
            e = enc.BaseEmbedder.pooling(types.SimpleNamespace(pooling_method=pm), hidden, m)
            out[f"pool/{pm}/{mname}/raw"] = e.numpy()
            out[f"pool/{pm}/{mname}/norm"] = norm(e).numpy()
        e = dec.BaseDecoderEmbedder.pooling(None, hidden, m)
        out[f"pool/last/{mname}/raw"] = e.numpy()
        out[f"pool/last/{mname}/norm"] = norm(e).numpy()
        e = mmeb.MMEBEmbeddingModel._pooling_last(None, hidden, m)
        out[f"pool/last_argmax/{mname}/raw"] = e.numpy()
        out[f"pool/last_argmax/{mname}/norm"] = norm(e).numpy()
        for pm in ["cls", "mean", "last_token"]:
            for nz in [True, False]:
                e = mm.MultimodalEmbedder._pool_hidden_states(
                    types.SimpleNamespace(pooling_method=pm, normalize_embeddings=nz), hidden, m)
                out[f"pool/mm_{pm}/{mname}/{'norm' if nz else 'raw'}"] = e.numpy()
    # half-precision inputs, as produced when use_fp16 is set (base.py:272-273)
    h16 = hidden.half()
    for pm in ["cls", "mean"]:
        e = enc.BaseEmbedder.pooling(types.SimpleNamespace(pooling_method=pm), h16.float() if pm == "x" else h16, right)
        out[f"pool16/{pm}/right/raw"] = e.float().numpy()
        out[f"pool16/{pm}/right/norm"] = norm(e.float()).numpy()
    # zero vector under F.normalize (eps keeps it 0) -- SURVEY §8(c)
    z = torch.zeros(2, 5)
    z[1, 2] = 3.0
    out["pool/zero_norm_in"] = z.numpy()
    out["pool/zero_norm_out"] = norm(z).numpy()

    path = os.path.join(HERE, "reference_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
