"""K2T's experimental mode (search_stream_tc.cu; off by default, behind NXF_K2T_FOLD) must give the same answers as the shipped path:
the threshold folded into the MMA (whole-row filter on score - tau), with each number of epilogue warp sets.  Short rows (one and two
K blocks), query counts on both sides of the 64-column accumulator stride, a corpus large enough for the sampled-threshold launches,
raw candidate pass and certified search."""
import os

import numpy as np
import pytest

from _gpu_common import run_and_check

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300, method="thread")]   # a hung kernel must not hold the box

VARIANTS = {
    "fold": {"NXF_K2T_FOLD": "1"},
    "fold_sets2": {"NXF_K2T_FOLD": "1", "NXF_K2T_SETS": "2"},
    "fold_sets4": {"NXF_K2T_FOLD": "1", "NXF_K2T_SETS": "4"},
}


@pytest.fixture
def k2t_env(request):
    saved = {k: os.environ.get(k) for k in ("NXF_K2T_FOLD", "NXF_K2T_SETS")}
    os.environ.update(VARIANTS[request.param])     # the library reads its switches with getenv at every call
    yield request.param
    for k, v in saved.items():
        if v is None:
            os.environ.pop(k, None)
        else:
            os.environ[k] = v


@pytest.mark.parametrize("k2t_env", sorted(VARIANTS), indirect=True)
@pytest.mark.parametrize("nq,dim,k", [(128, 64, 10), (48, 64, 100), (97, 128, 10), (33, 64, 1)])
def test_k2t_variant_matches_oracle(k2t_env, nq, dim, k):
    rng = np.random.default_rng(1234 + nq + dim)
    n = 148 * 128 * 70 + 77        # >= 68 tiles per CTA: estimate + sample + main launches, ragged last tile
    c = rng.standard_normal((n, dim), dtype=np.float32)
    c /= np.linalg.norm(c, axis=1, keepdims=True)
    x = rng.standard_normal((nq, dim), dtype=np.float32)
    x /= np.linalg.norm(x, axis=1, keepdims=True)
    run_and_check(c, x, k, store="bf16", path="tstream", expect_path="tstream")
    run_and_check(c, x, k, store="bf16", path="tstream_raw", rtol=2e-4, atol=2e-5, expect_path="tstream")


@pytest.mark.parametrize("k2t_env", ["fold"], indirect=True)
def test_fold_unnormalised_and_f32_plane(k2t_env):
    """Unnormalised rows (thresholds far from 1, negative scores: the -inf stand-in is the Cauchy-Schwarz bound) and fp32 storage
    (K2T streams the bf16 plane, the certificate + re-score make it exact)."""
    rng = np.random.default_rng(77)
    n, dim = 148 * 128 * 20 + 5, 64
    c = (rng.standard_normal((n, dim)) * rng.uniform(0.1, 30.0, size=(n, 1))).astype(np.float32)
    x = (rng.standard_normal((64, dim)) * 7.0 - 1.0).astype(np.float32)
    run_and_check(c, x, 10, store="bf16", path="tstream", expect_path="tstream")
    run_and_check(c, x, 10, store="f32", path="auto")
