"""Shared test helpers: golden fixtures + the seeded input recipe (same as tests/golden/make_golden.py)."""
import hashlib
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
GOLDEN_PATH = os.path.join(HERE, "golden", "reference_golden.npz")
_golden = None


def golden():
    global _golden
    if _golden is None:
        _golden = np.load(GOLDEN_PATH)
    return _golden


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def gen_inputs(n, d, q, seed, normalize=True, dtype=np.float32):
    rc = np.random.default_rng(seed)
    rq = np.random.default_rng(seed + 1000)
    c = rc.standard_normal((n, d), dtype=np.float32)
    x = rq.standard_normal((q, d), dtype=np.float32)
    if normalize:
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    return c.astype(dtype), x.astype(dtype)


def search_case_names():
    return sorted({k.split("/")[1] for k in golden().files if k.startswith("search/")})


def load_search_case(name):
    """Returns (corpus, queries, k, ref_scores, ref_ids); inputs regenerated from the seed and
    checked against the sha256 recorded when the reference produced the outputs."""
    g = golden()
    n, d, q, k, seed, norm, is16 = [int(v) for v in g[f"search/{name}/meta"]]
    if f"search/{name}/corpus" in g.files:
        c, x = g[f"search/{name}/corpus"], g[f"search/{name}/queries"]
    else:
        c, x = gen_inputs(n, d, q, seed, bool(norm), np.float16 if is16 else np.float32)
    assert sha(c) == bytes(g[f"search/{name}/corpus_sha"]).hex(), f"{name}: regenerated corpus differs from the golden run"
    assert sha(x) == bytes(g[f"search/{name}/query_sha"]).hex(), f"{name}: regenerated queries differ from the golden run"
    return c, x, k, g[f"search/{name}/scores"], g[f"search/{name}/ids"]
