"""CPU: host-side logic (sharding arithmetic, argument parsing, faiss shim plumbing, file formats)."""
import os

import numpy as np
import pytest

from nexus_b200 import shard_bounds
from nexus_b200.flat import parse_device, parse_store_dtype, _as_rows
from nexus_b200 import _lib as L


def test_shard_bounds_cover_and_balance():
    for n in [0, 1, 7, 8, 9, 8_841_823, 21_000_000]:
        for g in [1, 2, 3, 4, 8]:
            b = shard_bounds(n, g)
            assert b[0][0] == 0 and b[-1][1] == n and len(b) == g
            assert all(b[i][1] == b[i + 1][0] for i in range(g - 1))
            sizes = [hi - lo for lo, hi in b]
            assert max(sizes) - min(sizes) <= 1
    assert shard_bounds(21_000_000, 8)[3] == (7_875_000, 10_500_000)     # config 3: 2 625 000 rows per GPU


def test_parse_device_and_dtype():
    assert parse_device(3) == 3 and parse_device("cuda:5") == 5 and parse_device("2") == 2
    with pytest.raises(ValueError):
        parse_device("cpu")
    assert parse_store_dtype("bf16") == L.BF16 and parse_store_dtype("float16") == L.F16 and parse_store_dtype("fp32") == L.F32
    import torch
    assert parse_store_dtype(torch.bfloat16) == L.BF16
    with pytest.raises(ValueError):
        parse_store_dtype("int8")


def test_row_views():
    import torch
    a = np.arange(12, dtype=np.float64).reshape(3, 4)
    b = _as_rows(a, 4, "t")
    assert b.dtype == L.F32 and b.ld == 4 and b.n == 3 and b.on_device == 0      # reference casts to fp32 (utils.py:169)
    b = _as_rows(a.astype(np.float16)[:, :], 4, "t")
    assert b.dtype == L.F16
    b = _as_rows(np.zeros(4, np.float32), 4, "t")
    assert b.n == 1
    t = torch.zeros(5, 8, dtype=torch.bfloat16)
    b = _as_rows(t[:, :4], None, "t")          # a strided HOST view is densified: the host paths copy n*ld contiguous elements,
    assert b.dtype == L.BF16 and b.ld == 4 and b.d == 4   # which would run past the storage on the last row
    b = _as_rows(t[:, 4:], None, "t")
    assert b.ld == 4 and b.keep.is_contiguous()
    with pytest.raises(ValueError):
        _as_rows(a, 5, "t")
    with pytest.raises(ValueError):
        _as_rows(np.zeros((2, 2, 2)), None, "t")


def test_faiss_shim_surface():
    from nexus_b200 import faiss_shim as F
    idx = F.index_factory(16, "Flat", F.METRIC_INNER_PRODUCT)
    assert idx.d == 16 and idx.ntotal == 0 and idx.is_trained
    idx.nprobe = 10                              # recommendation/base.py:117
    idx.train(np.zeros((2, 16), np.float32))     # utils.py:188: no-op
    assert F.IndexFlatL2(8).metric_type == F.METRIC_L2 and F.IndexFlatIP(8).metric_type == F.METRIC_INNER_PRODUCT
    co = F.GpuMultipleClonerOptions()
    co.shard = True
    co.useFloat16 = True
    F.StandardGpuResources()
    with pytest.raises(ValueError):
        F.index_factory(16, "HNSW32")
    F.index_factory(16, "IVF100,Flat")           # served exactly (Nexus/inference/utils.py:19)


def test_index_file_roundtrip_header(tmp_path):
    from nexus_b200 import faiss_shim as F
    p = tmp_path / "x.nxf"

    class Fake:
        d, metric_type, ntotal = 4, 0, 3
        def reconstruct_n(self):
            return np.arange(12, dtype=np.float32).reshape(3, 4)
    F.write_index(Fake(), str(p))
    raw = open(p, "rb").read()
    assert raw[:8] == b"NXFLAT01" and len(raw) == 8 + 24 + 48
    with pytest.raises(ValueError):
        bad = tmp_path / "bad.bin"
        bad.write_bytes(b"0123456789abcdef")
        F.read_index(str(bad))


def test_install_as_faiss_registers_modules():
    import sys
    from nexus_b200 import faiss_shim as F
    saved = {k: sys.modules.get(k) for k in ("faiss", "faiss.contrib", "faiss.contrib.torch_utils")}
    try:
        m = F.install_as_faiss(force=True)
        import faiss
        import faiss.contrib.torch_utils  # noqa: F401  (modeling.py:8-9)
        assert faiss is m and hasattr(faiss, "IndexFlatIP") and hasattr(faiss, "index_cpu_to_all_gpus")
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def _perm_tile(s, ns, s0, n0, v):
    """Python restatement of perm_tile() in nexus_b200/csrc/params.cuh (the sample-first order of the corpus tiles)."""
    if s <= 1:
        return v
    if v >= ns:
        j = v - ns
        return j + j // (s - 1) + 1
    i = v
    if s0 > 1:
        if v < n0:
            i = v * s0
        else:
            j = v - n0
            i = j + j // (s0 - 1) + 1
    return i * s


@pytest.mark.parametrize("n_tiles,s,s0", [(1024, 16, 16), (4297, 16, 16), (34375, 16, 16), (78125, 31, 17), (1000, 6, 1), (37, 4, 3), (2, 2, 2)])
def test_tile_permutation_is_a_bijection_with_spread_samples(n_tiles, s, s0):
    ns = -(-n_tiles // s)
    n0 = -(-ns // s0) if s0 > 1 else ns
    phys = [_perm_tile(s, ns, s0, n0, v) for v in range(n_tiles)]
    assert sorted(phys) == list(range(n_tiles))
    assert all(p % s == 0 for p in phys[:ns])                     # the sample: every s-th tile
    if s0 > 1:
        assert all(p % (s * s0) == 0 for p in phys[:n0])          # its first level: every (s*s0)-th tile
    assert phys[ns:] == sorted(phys[ns:])                         # everything else in corpus order


def test_load_embedding_cache_formats(tmp_path):
    """The cache files the reference writes: .npy, item_vectors.pt, MMEB pickles -- host parsing only."""
    import pickle
    import torch
    from nexus_b200 import faiss_shim as F
    rng = np.random.default_rng(0)
    rows = rng.standard_normal((7, 16)).astype(np.float32)
    np.save(tmp_path / "doc.npy", rows)                                                   # searcher.py:126-127
    r, ids = F.load_embedding_cache(str(tmp_path / "doc.npy"))
    assert ids is None and np.array_equal(np.asarray(r), rows)
    torch.save({"item_vectors": torch.from_numpy(rows), "item_ids": torch.arange(100, 107)}, tmp_path / "item_vectors.pt")   # trainer.py:38-39
    r, ids = F.load_embedding_cache(str(tmp_path / "item_vectors.pt"))
    assert np.array_equal(r, rows) and ids.tolist() == list(range(100, 107))
    with open(tmp_path / "qry.pkl", "wb") as f:
        pickle.dump(rows, f)                                                              # eval_embedding.py:269-270
    r, ids = F.load_embedding_cache(str(tmp_path / "qry.pkl"))
    assert ids is None and np.array_equal(r, rows)
    d = {f"c{i}": rows[i] for i in (3, 1, 5)}
    with open(tmp_path / "cand.pkl", "wb") as f:
        pickle.dump(d, f)                                                                 # eval_embedding.py:309-310
    r, ids = F.load_embedding_cache(str(tmp_path / "cand.pkl"))
    assert ids == ["c3", "c1", "c5"] and np.array_equal(r, rows[[3, 1, 5]])               # dictionary order, :353-354
    torch.save({"something": 1}, tmp_path / "other.pt")
    with pytest.raises(ValueError):
        F.load_embedding_cache(str(tmp_path / "other.pt"))


def test_ivfflat_constructor_shape():
    """Nexus/inference/utils.py:19-20: IndexFlatIP quantizer + IndexIVFFlat(quantizer, d, 100, METRIC_INNER_PRODUCT); no device
    is touched until the first add."""
    from nexus_b200 import faiss_shim as F
    quantizer = F.IndexFlatIP(32)
    index = F.IndexIVFFlat(quantizer, 32, 100, F.METRIC_INNER_PRODUCT)
    assert index.d == 32 and index.nlist == 100 and index.ntotal == 0 and index.is_trained
    index.nprobe = 10                      # recommendation/base.py:117: settable, ignored
    index.train(np.zeros((4, 32), np.float32))
    assert F.IndexIVFFlat(quantizer, 32, 100).metric_type == F.METRIC_L2   # Faiss's default metric for this class


def test_faiss_native_flat_file_layout(tmp_path):
    """Faiss's own flat file layout (index_write.cpp, restated; unpinned -- no Faiss in this image): header parse, payload,
    truncation and inconsistent headers are loud.  No device is touched: the parser is called directly."""
    import struct
    from nexus_b200 import faiss_shim as F
    rows = np.arange(15, dtype=np.float32).reshape(5, 3)

    class Fake:
        d, metric_type, ntotal = 3, F.METRIC_INNER_PRODUCT, 5
        def reconstruct_n(self):
            return rows
    p = tmp_path / "flat.index"
    F.write_index_faiss_flat(Fake(), str(p))
    raw = open(p, "rb").read()
    assert raw[:4] == b"IxFI" and len(raw) == 4 + 4 + 24 + 1 + 4 + 8 + 60
    with open(p, "rb") as f:
        f.seek(4)
        d, metric, got = F._read_faiss_flat(f, b"IxFI")
    assert (d, metric) == (3, F.METRIC_INNER_PRODUCT) and np.array_equal(got, rows)
    bad = tmp_path / "bad.index"                       # payload count does not match ntotal * d
    bad.write_bytes(raw[:37] + struct.pack("<Q", 14) + raw[45:])
    with open(bad, "rb") as f:
        f.seek(4)
        with pytest.raises(ValueError):
            F._read_faiss_flat(f, b"IxFI")
    trunc = tmp_path / "trunc.index"
    trunc.write_bytes(raw[:-8])
    with open(trunc, "rb") as f:
        f.seek(4)
        with pytest.raises(ValueError):
            F._read_faiss_flat(f, b"IxFI")


FAISS_NAMES_USED_BY_THE_REFERENCE = [   # grep -o "faiss\.\w+" over Nexus/**/*.py at the surveyed revision
    "read_index", "write_index", "IndexFlatIP", "IndexFlatL2", "IndexIVFFlat", "index_factory", "index_cpu_to_gpu", "index_cpu_to_all_gpus",
    "StandardGpuResources", "GpuMultipleClonerOptions", "METRIC_INNER_PRODUCT", "contrib",
]


def test_shim_covers_every_faiss_name_the_reference_uses():
    """Object-level drop-in: every `faiss.<name>` the reference touches exists on the shim once it is installed as
    `faiss`.  When the reference tree is present (build container), the committed list is re-derived from it."""
    import re
    import sys
    from nexus_b200 import faiss_shim as F
    saved = {k: sys.modules.get(k) for k in ("faiss", "faiss.contrib", "faiss.contrib.torch_utils")}
    try:
        m = F.install_as_faiss(force=True)
        for name in FAISS_NAMES_USED_BY_THE_REFERENCE:
            assert hasattr(m, name), name
        import faiss.contrib.torch_utils  # noqa: F401  (modeling.py:9)
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    ref = "/root/reference/Nexus"
    if os.path.isdir(ref):
        found = set()
        for dp, _, fs in os.walk(ref):
            for fn in fs:
                if fn.endswith(".py"):
                    found |= set(re.findall(r"\bfaiss\.(\w+)", open(os.path.join(dp, fn), encoding="utf-8", errors="ignore").read()))
        assert found <= set(FAISS_NAMES_USED_BY_THE_REFERENCE), sorted(found - set(FAISS_NAMES_USED_BY_THE_REFERENCE))
