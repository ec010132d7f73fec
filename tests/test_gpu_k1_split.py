"""K1's split accumulators (search_tc.cu, template SPLIT; rows of one or two K blocks with two epilogue warp sets): each 256-column
accumulator is computed and handed over as two 128-column halves, and under cta_group::2 a half holds corpus rows of BOTH CTAs'
slices, so the epilogue remaps slice -> corpus row; with four groups (the default on large short-row problems) each group waits for,
reads and hands back one half.  The planner only takes more than one set on large problems; NXF_K1_SETS forces them here
so that small cases cover the remapping: both CTA groupings, both metrics, a ragged last corpus tile, raw candidate pass and
certified search, fp16 and bf16 rows.  NXF_K1_SPLIT=2 makes the library fail the call if the split kernel cannot be used (so the
test knows it ran); NXF_K1_SPLIT=0 is the control (same answers from whole accumulators)."""
import os

import numpy as np
import pytest

from _gpu_common import run_and_check

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300, method="thread")]


@pytest.fixture
def k1_env(request):
    """request.param = "<sets>:<split>": NXF_K1_SETS (2 = two sets, each reading both halves of its buffer; 4 = four groups, one half
    each) and NXF_K1_SPLIT (2 = required, 0 = whole accumulators)."""
    keys = ("NXF_K1_SETS", "NXF_K1_SPLIT")
    saved = {k: os.environ.get(k) for k in keys}
    os.environ["NXF_K1_SETS"], os.environ["NXF_K1_SPLIT"] = request.param.split(":")
    yield request.param
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


@pytest.mark.parametrize("k1_env", ["4:2", "2:2", "2:0"], indirect=True)
@pytest.mark.parametrize("path", ["tensor", "tensor_1cta", "tensor_raw"])
@pytest.mark.parametrize("dim,store,metric", [(64, "bf16", "ip"), (128, "f16", "ip"), (100, "bf16", "l2"), (64, "bf16", "l2")])
def test_k1_split_matches_oracle(k1_env, path, dim, store, metric):
    if metric == "l2" and path == "tensor_raw":
        pytest.skip("the raw candidate pass is compared for the inner product only")
    rng = np.random.default_rng(4321 + dim)
    n, nq, k = 256 * 37 + 91, 300, 20          # 38 corpus tiles, the last one ragged; two query tiles (one partly filled)
    c = rng.standard_normal((n, dim), dtype=np.float32)
    x = rng.standard_normal((nq, dim), dtype=np.float32)
    if metric == "ip":
        c /= np.linalg.norm(c, axis=1, keepdims=True)
        x /= np.linalg.norm(x, axis=1, keepdims=True)
    raw = path == "tensor_raw"
    run_and_check(c, x, k, store=store, path=path, metric=metric, expect_path="tensor",
                  rtol=2e-4 if raw else 1e-5, atol=2e-5 if raw else 1e-6)


@pytest.mark.parametrize("k1_env", ["4:2", "2:2"], indirect=True)
def test_k1_split_ids_cover_every_slice(k1_env):
    """One planted best row per 32-row slice of a tile: every slice's remapped row index is exercised by an exact id check."""
    dim, n = 64, 256 * 8
    rng = np.random.default_rng(9)
    c = (0.01 * rng.standard_normal((n, dim))).astype(np.float32)
    x = np.zeros((8 * 8, dim), dtype=np.float32)
    want = []
    for t in range(8):                 # tile t
        for sl in range(8):            # slice sl of the tile: plant row t*256 + sl*32 + (t + sl) % 32 as query (t, sl)'s best
            r = t * 256 + sl * 32 + (t + sl) % 32
            q = t * 8 + sl
            x[q, (q * 7) % dim] = 1.0
            c[r, (q * 7) % dim] = 2.0 + 0.01 * q
            want.append(r)
    from nexus_b200 import FlatIndex
    from nexus_b200 import _lib as L
    idx = FlatIndex(dim, "ip", "bf16")
    idx.add(c)
    D, I = idx.search(x, 1, flags=L.SEARCH_FORCE_TENSOR)
    assert idx.stats()["last_path_name"] == "tensor"
    assert I[:, 0].tolist() == want
