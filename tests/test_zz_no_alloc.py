"""The allocation-free entry point nxf_pool_mean_normalize_static (include/nexus_flat.h; no counterpart in the reference)
against the existing nxf_pool_normalize(NXF_POOL_MEAN), both through the C ABI (ctypes), on two
calls in a row that share the library's static scratch array.  Collected last on purpose (file name):
a failure here must not hide other tests.  Passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _both(lib, L, torch, h, m, normalize, out_dtype):
    codes = {torch.float32: L.F32, torch.float16: L.F16, torch.bfloat16: L.BF16}
    B, Lq, H = h.shape
    dev = h.device.index
    stream = int(torch.cuda.current_stream(dev).cuda_stream)
    ref = torch.empty((B, H), dtype=out_dtype, device=h.device)
    got = torch.full((B, H), float("nan"), dtype=out_dtype, device=h.device)
    L.check(lib.nxf_pool_normalize(h.data_ptr(), codes[h.dtype], m.data_ptr(), B, Lq, H, L.POOL_MEAN, 1 if normalize else 0, 1e-12,
                                   ref.data_ptr(), codes[out_dtype], dev, stream))
    L.check(lib.nxf_pool_mean_normalize_static(h.data_ptr(), codes[h.dtype], m.data_ptr(), B, Lq, H, 1 if normalize else 0, 1e-12,
                                               got.data_ptr(), codes[out_dtype], dev, stream))
    torch.cuda.synchronize()
    return ref.float().cpu().numpy(), got.float().cpu().numpy()


def test_static_scratch_mean_pooling_equals_the_allocating_entry_point_twice_in_a_row():
    import torch
    from nexus_b200 import _lib as L
    lib = L.load()
    gen = torch.Generator().manual_seed(123)
    # two calls in a row on different shapes: the second finds the first one's partial sums and control words in the shared array
    for (B, Lq, H, dt, out_dt, normalize) in [(256, 512, 768, torch.float16, torch.float32, True),
                                              (37, 100, 384, torch.bfloat16, torch.float32, True),
                                              (256, 512, 768, torch.float16, torch.float16, False),
                                              (5, 9, 64, torch.float32, torch.float32, True)]:
        h = torch.randn(B, Lq, H, generator=gen).to(dt).cuda()
        lens = torch.randint(1, Lq + 1, (B,), generator=gen)
        m = (torch.arange(Lq)[None, :] < lens[:, None]).long().cuda()
        for _ in range(2):
            ref, got = _both(lib, L, torch, h, m, normalize, out_dt)
            assert np.isfinite(got).all()
            # same kernel, possibly another number of slices per row (fp32 summation order): the tolerance of tests/test_gpu_pool.py
            tol = dict(rtol=2e-5, atol=2e-6) if out_dt == torch.float32 else dict(rtol=2e-3, atol=2e-3)
            np.testing.assert_allclose(got, ref, **tol)


def test_static_scratch_rejects_sizes_beyond_its_maximum():
    import torch
    from nexus_b200 import _lib as L
    lib = L.load()
    h = torch.zeros(8, 4, 64, dtype=torch.float16, device="cuda")
    m = torch.ones(8, 4, dtype=torch.int64, device="cuda")
    out = torch.zeros(8, 64, dtype=torch.float32, device="cuda")
    stream = int(torch.cuda.current_stream().cuda_stream)
    # more rows than the control words of the static array hold (the pointers are never dereferenced: rejected on the host)
    rc = lib.nxf_pool_mean_normalize_static(h.data_ptr(), L.F16, m.data_ptr(), 4097, 4, 64, 1, 1e-12, out.data_ptr(), L.F32, 0, stream)
    assert rc == L.NXF_EINVAL and b"4096" in lib.nxf_last_error()
    # partial sums beyond 8 MB even with one slice per row: 4096 rows x 4096 columns x 4 bytes = 64 MB
    rc = lib.nxf_pool_mean_normalize_static(h.data_ptr(), L.F16, m.data_ptr(), 4096, 4, 4096, 1, 1e-12, out.data_ptr(), L.F32, 0, stream)
    assert rc == L.NXF_EINVAL and b"static scratch" in lib.nxf_last_error()
    # a row length the vectorised kernel does not take
    rc = lib.nxf_pool_mean_normalize_static(h.data_ptr(), L.F16, m.data_ptr(), 8, 4, 63, 1, 1e-12, out.data_ptr(), L.F32, 0, stream)
    assert rc == L.NXF_EUNSUPPORTED
    with pytest.raises(ValueError):
        L.check(lib.nxf_pool_mean_normalize_static(h.data_ptr(), L.F16, 0, 8, 4, 64, 1, 1e-12, out.data_ptr(), L.F32, 0, stream))
    torch.cuda.synchronize()
