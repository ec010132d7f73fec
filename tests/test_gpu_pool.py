"""GPU parity of the fused pool + normalise kernel (K0) against the reference's golden outputs and the oracle."""
import numpy as np
import pytest

import oracle
from _common import golden

pytestmark = pytest.mark.gpu

POOL_MAP = {"cls": "cls", "mean": "mean", "last": "last", "last_argmax": "last_argmax"}


@pytest.mark.parametrize("mask", ["right", "left", "holes", "full"])
@pytest.mark.parametrize("kind", sorted(POOL_MAP))
def test_pool_matches_reference_golden(kind, mask):
    import torch
    from nexus_b200 import pool_and_normalize
    g = golden()
    h = torch.from_numpy(g["pool/hidden"]).cuda()
    m = torch.from_numpy(g[f"pool/mask_{mask}"]).cuda()
    raw = pool_and_normalize(h, m, POOL_MAP[kind], normalize=False)
    nrm = pool_and_normalize(h, m, POOL_MAP[kind], normalize=True)
    np.testing.assert_allclose(raw.cpu().numpy(), g[f"pool/{kind}/{mask}/raw"], rtol=1e-5, atol=1e-6)
    np.testing.assert_allclose(nrm.cpu().numpy(), g[f"pool/{kind}/{mask}/norm"], rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("dtype", ["float32", "float16", "bfloat16"])
@pytest.mark.parametrize("shape", [(7, 33, 384), (16, 128, 768), (3, 17, 1024), (5, 9, 100), (2, 512, 1536)])
@pytest.mark.parametrize("pm", ["cls", "mean", "last", "last_argmax"])
def test_pool_vs_oracle(pm, shape, dtype):
    import torch
    from nexus_b200 import pool_and_normalize
    B, L, H = shape
    gen = torch.Generator().manual_seed(B * 1000 + L)
    h = torch.randn(B, L, H, generator=gen).to(getattr(torch, dtype))
    lens = torch.randint(1, L + 1, (B,), generator=gen)
    m = (torch.arange(L)[None, :] < lens[:, None]).long()
    out = pool_and_normalize(h.cuda(), m.cuda(), pm, normalize=True, out_dtype=torch.float32)
    ref = oracle.pool_and_normalize(h.float().numpy(), m.numpy(), pm, normalize=True)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=2e-5, atol=2e-6)
    # cast-on-write variant (straight into a bf16 shard)
    out16 = pool_and_normalize(h.cuda(), m.cuda(), pm, normalize=True, out_dtype=torch.bfloat16)
    np.testing.assert_allclose(out16.float().cpu().numpy(), ref, rtol=1e-2, atol=1e-3)


def test_pool_zero_vector_and_numpy_form():
    import torch
    from nexus_b200 import pool_and_normalize
    g = golden()
    z = torch.from_numpy(g["pool/zero_norm_in"]).cuda()[:, None, :]      # [2, 1, 5], cls pooling
    out = pool_and_normalize(z, None, "cls", normalize=True)
    np.testing.assert_array_equal(out.cpu().numpy(), g["pool/zero_norm_out"])
    out2 = pool_and_normalize(z, None, "cls", normalize=True, eps=0.0)    # base.py:673: no clamp -> NaN
    assert torch.isnan(out2[0]).all() and not torch.isnan(out2[1]).any()


def test_pool_unknown_method_and_cpu_tensor():
    import torch
    from nexus_b200 import pool_and_normalize
    with pytest.raises(NotImplementedError):
        pool_and_normalize(torch.zeros(1, 2, 8).cuda(), torch.ones(1, 2).cuda(), "max")
    with pytest.raises(RuntimeError):
        pool_and_normalize(torch.zeros(1, 2, 8), torch.ones(1, 2), "cls")


def test_encode_to_index_handoff_stays_on_device():
    """N1: pooled + normalised bf16 embeddings go straight into the index, no host round trip."""
    import torch
    from nexus_b200 import FlatIndex, pool_and_normalize
    gen = torch.Generator().manual_seed(3)
    h = torch.randn(500, 12, 128, generator=gen).half().cuda()
    m = torch.ones(500, 12, dtype=torch.long).cuda()
    emb = pool_and_normalize(h, m, "mean", normalize=True, out_dtype=torch.bfloat16)
    idx = FlatIndex(128, "ip", "bf16")
    idx.add(emb)
    D, I = idx.search(emb[:64], 1)
    assert (I[:, 0].cpu() == torch.arange(64)).all()                       # every row's nearest neighbour is itself
    assert torch.allclose(D[:, 0].cpu(), torch.ones(64), atol=2e-2)


@pytest.mark.parametrize("cache", ["0", "1"])
def test_mean_pool_workspace_reuse_across_shapes_and_streams(cache, monkeypatch):
    """NXF_POOL_WS_CACHE=1: the split mean kernel's workspace is cached per (device, stream) and its control words must be back at
    zero after every launch: repeated calls, growing / shrinking batch sizes, and a second stream all give the oracle's result
    (=0: one stream-ordered allocation per call)."""
    import torch
    monkeypatch.setenv("NXF_POOL_WS_CACHE", cache)
    import oracle
    from nexus_b200 import pool_and_normalize
    gen = torch.Generator().manual_seed(9)
    side = torch.cuda.Stream()
    for rep, (B, Lq, H) in enumerate([(4, 256, 384), (64, 512, 768), (3, 1024, 128), (64, 512, 768), (130, 64, 256), (4, 256, 384)]):
        h = torch.randn(B, Lq, H, generator=gen).half()
        lens = torch.randint(1, Lq + 1, (B, 1), generator=gen)
        m = (torch.arange(Lq)[None, :] < lens).long()
        ref = oracle.pool_and_normalize(h.float().numpy(), m.numpy(), "mean", normalize=True)
        hd, md = h.cuda(), m.cuda()
        for _ in range(3):
            out = pool_and_normalize(hd, md, "mean", normalize=True)
            np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=2e-5, atol=2e-6)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            out2 = pool_and_normalize(hd, md, "mean", normalize=rep % 2 == 0)
        side.synchronize()
        ref2 = ref if rep % 2 == 0 else oracle.pool_and_normalize(h.float().numpy(), m.numpy(), "mean", normalize=False)
        np.testing.assert_allclose(out2.cpu().numpy(), ref2, rtol=2e-5, atol=2e-6)
