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


# ---- MMEB ranking cases (tests/golden/make_golden_mmeb.py runs the reference's lines on exactly these inputs) ----
MMEB_GOLDEN_PATH = os.path.join(HERE, "golden", "mmeb_golden.npz")
MMEB_CASES = {
    # name: (seed, Nq, Nc, D, longest list, duplicates inside lists)
    "small": (31, 40, 300, 48, 60, False),
    "wide": (32, 25, 500, 384, 200, False),
    "dups": (33, 30, 120, 64, 40, True),
}


def gen_mmeb_case(name):
    """-> (qry_embeds [Nq, D] f32, cand_embed_dict {name: [D] f32}, gt_infos) -- the shapes eval_embedding.py:337-345 loads."""
    seed, nq, nc, d, longest, dups = MMEB_CASES[name]
    rng = np.random.default_rng(seed)
    cand = rng.standard_normal((nc, d), dtype=np.float32)
    cand /= np.linalg.norm(cand, axis=1, keepdims=True)
    noise = rng.standard_normal((nq, d), dtype=np.float32)
    noise /= np.linalg.norm(noise, axis=1, keepdims=True)
    names = [f"cand_{i:05d}" for i in rng.permutation(nc)]           # dictionary order is not the numeric order
    cand_embed_dict = {n: cand[i] for i, n in enumerate(names)}
    gt_infos = []
    for qi in range(nq):
        ln = int(rng.integers(1, longest + 1))
        pick = rng.choice(nc, size=ln, replace=dups)
        cand_names = [names[int(j)] for j in pick]
        if qi % 3 == 0:
            label = cand_names[int(rng.integers(0, ln))]               # a single string (eval_embedding.py:371)
        else:
            label = [cand_names[int(j)] for j in rng.choice(ln, size=min(ln, 2), replace=False)]
        gt_infos.append({"label_name": label, "cand_names": cand_names})
    # a query looks like its (first) label plus noise, so the metrics are neither 0 nor 1
    first = [g["label_name"] if isinstance(g["label_name"], str) else g["label_name"][0] for g in gt_infos]
    qry = np.stack([cand_embed_dict[n] for n in first]) + np.float32(3.0) * noise
    qry /= np.linalg.norm(qry, axis=1, keepdims=True)
    return qry.astype(np.float32), cand_embed_dict, gt_infos


_mmeb_golden = None


def mmeb_golden():
    global _mmeb_golden
    if _mmeb_golden is None:
        _mmeb_golden = np.load(MMEB_GOLDEN_PATH)
    return _mmeb_golden


def load_mmeb_case(name):
    """Inputs regenerated from the seed and checked against the sha256 stored when the reference produced the outputs."""
    g = mmeb_golden()
    qry, cdict, gts = gen_mmeb_case(name)
    assert sha(qry) == bytes(g[f"{name}/qry_sha"]).hex(), f"{name}: regenerated queries differ from the golden run"
    assert sha(np.stack(list(cdict.values()))) == bytes(g[f"{name}/cand_sha"]).hex(), f"{name}: regenerated candidates differ"
    return qry, cdict, gts


def ranking_agrees(ref_scores, ref_ranked, got_ranked, rtol=1e-5, atol=1e-6):
    """Two rankings of the same list agree if at every rank the scores (looked up in ref_scores) match within tolerance:
    positions may only differ among scores tied within tolerance."""
    ref_ranked = np.asarray(ref_ranked, dtype=np.int64)
    got_ranked = np.asarray(got_ranked, dtype=np.int64)
    if ref_ranked.shape != got_ranked.shape:
        return False
    a, b = ref_scores[ref_ranked], ref_scores[got_ranked]
    return bool(np.all(np.abs(a - b) <= atol + rtol * np.abs(a)))
