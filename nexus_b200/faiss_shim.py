"""A minimal ``faiss``-shaped module backed by the B200 engine.

Covers exactly the Faiss names the reference's hot path touches:
  * ``index_factory(d, "Flat", METRIC_INNER_PRODUCT)``          Nexus/evaluation/text_retrieval/utils.py:176
  * ``GpuMultipleClonerOptions`` / ``index_cpu_to_all_gpus``      utils.py:180-183
  * ``IndexFlatIP(d)`` / ``IndexFlatL2(d)``                       Nexus/training/embedder/recommendation/modeling.py:184-190
  * ``StandardGpuResources()`` / ``index_cpu_to_gpu(res, dev, index)``   modeling.py:182,194; recommendation/base.py:115-116
  * ``read_index`` / ``write_index``                             serve/inference_service_text_retrieval.py:35; Nexus/inference/utils.py:25
  * ``faiss.contrib.torch_utils`` (import-only; torch tensors are accepted natively)   modeling.py:9

Everything is exact flat search; ``IVF*`` factory strings (Nexus/inference/utils.py:19) are served by an
exact index as well -- a recall-superset of IVFFlat(nprobe<nlist) -- and ``nprobe`` is accepted and ignored.
Objects created "on the CPU" (``IndexFlatIP(d)`` before ``index_cpu_to_gpu``) only record their parameters and
buffered rows; compute always happens on the GPU: there is no CPU search path.
"""
from __future__ import annotations

import os
import re
import sys
import types
from typing import Optional

import numpy as np

from .flat import FlatIndex, _is_tensor
from .sharded import ShardedFlatIndex

METRIC_INNER_PRODUCT = 0
METRIC_L2 = 1


def _store_dtype_default() -> str:
    return os.environ.get("NEXUS_B200_STORE_DTYPE", "f32")


class _ItemCache:
    """Row N3: ``BaseRetriever.eval_step`` builds ``IndexFlatIP(d)`` -> ``index_cpu_to_gpu`` -> ``add(item_vectors)``
    again for EVERY evaluation batch (modeling.py:182-195, called per batch from evaluation/recommendation/evaluator.py:115-134)
    with the same item matrix.  The shim keeps the device shard it built and hands it to the next identical ``add``:

    * same tensor OBJECT (a weak reference is still alive, so its storage cannot have been recycled), same ``data_ptr`` and same
      ``_version`` -> same contents, no device work at all (Inner-product / Euclidean scorers pass ``self.item_vectors`` itself);
    * otherwise (the Cosine scorer normalises into a NEW tensor each batch, modeling.py:188) a checksum of the rows decides: two
      device reductions over the tensor's bits and one 16-byte read-back instead of a convert + copy of all N rows.

    A few entries per process (``NEXUS_B200_SHIM_CACHE`` = entries, 0 disables).  A cached shard is shared read-only: a further
    ``add`` / ``reset`` on an object that holds one first takes a private copy."""

    def __init__(self):
        self.entries = []       # dicts: key, ref, ptr, version, checksum, impl

    @staticmethod
    def capacity() -> int:
        try:
            return max(0, int(os.environ.get("NEXUS_B200_SHIM_CACHE", "2")))
        except ValueError:
            return 2

    @staticmethod
    def checksum(x):
        import torch
        v = x.contiguous().view(-1).view(torch.int32 if x.element_size() == 4 else torch.int16)
        both = torch.stack([v.sum(dtype=torch.int64), v[::3].sum(dtype=torch.int64)])
        return tuple(both.tolist())

    def lookup(self, x, key):
        for e in self.entries:
            if e["key"] != key:
                continue
            if e["ref"]() is x and e["ptr"] == x.data_ptr() and e["version"] == x._version:
                return e["impl"], "identity"
        cs = None
        for e in self.entries:
            if e["key"] != key:
                continue
            cs = cs or self.checksum(x)
            if e["checksum"] == cs:
                return e["impl"], "checksum"
        return None, cs

    def store(self, x, key, impl, cs=None):
        import weakref
        cap = self.capacity()
        if cap == 0:
            return
        self.entries.insert(0, {"key": key, "ref": weakref.ref(x), "ptr": x.data_ptr(), "version": x._version,
                                "checksum": cs or self.checksum(x), "impl": impl})
        del self.entries[cap:]

    def clear(self):
        self.entries.clear()


_item_cache = _ItemCache()
cache_stats = {"identity": 0, "checksum": 0, "miss": 0}


class _LazyFlat:
    """IndexFlat{IP,L2}(d): materialises a device index on first use (device chosen by
    index_cpu_to_gpu / index_cpu_to_all_gpus, else the current device)."""
    __nexus_b200_lazy__ = True

    def __init__(self, d: int, metric: int, store_dtype: Optional[str] = None):
        self.d = int(d)
        self.metric_type = metric
        self.nprobe = 1
        self.is_trained = True
        self._store_dtype = store_dtype or _store_dtype_default()
        self._impl = None
        self._devices = None      # None -> current device; list -> sharded
        self._shared = False      # _impl is also held by the item cache: copy before mutating

    def _ensure(self, like=None):
        if self._impl is None:
            metric = "l2" if self.metric_type == METRIC_L2 else "ip"
            if self._devices is not None and len(self._devices) > 1:
                self._impl = ShardedFlatIndex(self.d, metric, self._store_dtype, devices=self._devices)
            else:
                dev = self._devices[0] if self._devices else (like.device.index if (like is not None and _is_tensor(like) and like.is_cuda) else None)
                self._impl = FlatIndex(self.d, metric, self._store_dtype, device=dev)
        return self._impl

    @property
    def ntotal(self) -> int:
        return 0 if self._impl is None else self._impl.ntotal

    def train(self, x=None):
        return None

    def _own(self):
        """Take a private copy of a shard that the item cache also holds (device-to-device, fp32 round trip of the stored values)."""
        if self._shared and self._impl is not None:
            import torch
            old = self._impl
            rows = torch.empty((old.ntotal, old.d), dtype=torch.float32, device=f"cuda:{old.device}")
            old.reconstruct_into(rows)
            self._impl = FlatIndex(self.d, old.metric, old.store_dtype, device=old.device)
            if old.ntotal:
                self._impl.add(rows)
        self._shared = False

    def add(self, x):
        single = self._devices is None or len(self._devices) == 1
        if self._impl is None and single and _is_tensor(x) and x.is_cuda and x.dim() == 2 and _ItemCache.capacity() > 0:
            dev = self._devices[0] if self._devices else x.device.index
            if dev == x.device.index:
                key = (dev, self.d, int(self.metric_type), self._store_dtype, tuple(x.shape), str(x.dtype))
                impl, how = _item_cache.lookup(x, key)
                if impl is not None:
                    cache_stats[how] += 1
                    self._impl, self._shared = impl, True
                    return
                cache_stats["miss"] += 1
                self._ensure(x).add(x)
                _item_cache.store(x, key, self._impl, how)
                self._shared = True
                return
        self._own()
        self._ensure(x).add(x)

    def search(self, x, k):
        return self._ensure(x).search(x, k)

    def reset(self):
        if self._shared:
            self._impl, self._shared = None, False
        elif self._impl is not None:
            self._impl.reset()

    def reconstruct_n(self, first=0, n=None):
        impl = self._ensure()
        if isinstance(impl, ShardedFlatIndex):
            return np.concatenate([s.reconstruct_n() for s in impl.shards if s.ntotal], axis=0)[first:None if n is None else first + n]
        return impl.reconstruct_n(first, n)


class IndexFlatIP(_LazyFlat):
    def __init__(self, d: int):
        super().__init__(d, METRIC_INNER_PRODUCT)


class IndexFlatL2(_LazyFlat):
    def __init__(self, d: int):
        super().__init__(d, METRIC_L2)


class IndexIVFFlat(_LazyFlat):
    """``faiss.IndexIVFFlat(quantizer, d, nlist, metric)`` as the recsys index builder constructs it
    (Nexus/inference/utils.py:19-22: ``IndexFlatIP`` quantizer, nlist = 100, ``METRIC_INNER_PRODUCT``), served by the exact
    flat scan: every list is probed, so the result is what IVF returns at ``nprobe = nlist`` -- a recall superset of any
    smaller ``nprobe`` (settable and ignored, recommendation/base.py:117).  Faiss's default metric for this class is L2."""

    def __init__(self, quantizer, d: int, nlist: int = 100, metric: int = METRIC_L2):
        super().__init__(int(d), int(metric))
        self.quantizer, self.nlist = quantizer, int(nlist)


def index_factory(d: int, description: str = "Flat", metric: int = METRIC_INNER_PRODUCT):
    desc = description.strip()
    if desc != "Flat" and not re.match(r"^(IVF\d+,)?Flat$", desc):
        raise ValueError(f"index_factory({description!r}): only exact flat indices are implemented")
    return _LazyFlat(d, metric)


class StandardGpuResources:
    def noTempMemory(self):
        return None

    def setTempMemory(self, _bytes):
        return None


class GpuMultipleClonerOptions:
    def __init__(self):
        self.shard = False
        self.useFloat16 = False


class GpuClonerOptions(GpuMultipleClonerOptions):
    pass


def get_num_gpus() -> int:
    import torch
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def index_cpu_to_gpu(res, device: int, index, options=None):
    """``faiss.index_cpu_to_gpu(res, device, index)`` (modeling.py:194 passes ``index=`` by keyword; recommendation/base.py:115-116
    moves a populated index read from disk).  An empty index only records its device.  A populated one that already lives on
    ``device`` with the requested storage type is returned as it is (no second upload); otherwise its rows move device to device
    (or host to device for a sharded source) into a new shard.  ``options.useFloat16`` selects fp16 storage in both cases."""
    device = int(device)
    want16 = options is not None and getattr(options, "useFloat16", False)
    store = "f16" if want16 else index._store_dtype
    impl = index._impl
    if impl is None or index.ntotal == 0:
        if impl is not None and not (isinstance(impl, FlatIndex) and impl.device == device and impl.store_dtype == store):
            index._impl, index._shared = None, False     # an emptied shard on another device / of another type: rebuilt on first use
        index._devices = [device]
        index._store_dtype = store
        return index
    if isinstance(impl, FlatIndex) and impl.device == device and impl.store_dtype == store:
        index._devices = [device]
        return index
    new = _LazyFlat(index.d, index.metric_type, store)
    new._devices = [device]
    if isinstance(impl, FlatIndex):
        import torch
        rows = torch.empty((impl.ntotal, impl.d), dtype=torch.float32, device=f"cuda:{impl.device}")
        impl.reconstruct_into(rows)
        new._ensure().add(rows.to(f"cuda:{device}"))
    else:
        new._ensure().add(index.reconstruct_n())
    return new


def index_cpu_to_all_gpus(index, co=None, ngpu: int = -1):
    n = get_num_gpus() if ngpu < 0 else ngpu
    if n == 0:
        raise RuntimeError("index_cpu_to_all_gpus: no CUDA device (there is no CPU fallback)")
    if co is not None and getattr(co, "useFloat16", False):
        index._store_dtype = "f16"         # utils.py:182
    rows = index.reconstruct_n() if index.ntotal else None
    new = _LazyFlat(index.d, index.metric_type, index._store_dtype)
    new._devices = list(range(n))          # row shards (co.shard = True, utils.py:181); replicas are not implemented
    if rows is not None:
        new.add(rows)
    return new


def index_gpu_to_cpu(index):
    return index


_MAGIC = b"NXFLAT01"


def write_index(index, path: str) -> None:
    """Self-describing flat file: magic, d, metric, n, then fp32 rows (np.save-compatible payload)."""
    rows = index.reconstruct_n() if index.ntotal else np.empty((0, index.d), np.float32)
    with open(path, "wb") as f:
        f.write(_MAGIC)
        np.array([index.d, index.metric_type, rows.shape[0]], dtype=np.int64).tofile(f)
        rows.astype(np.float32).tofile(f)


def _read_faiss_flat(f, fourcc: bytes):
    """Native Faiss flat index file ('IxFI' = IndexFlatIP, 'IxF2' = IndexFlatL2, 'IxFl' = old IndexFlat), restated from Faiss's
    published writer (faiss/impl/index_write.cpp: write_index_header + the flat payload): int32 d, int64 ntotal, 2 x int64 dummy,
    uint8 is_trained, int32 metric_type (+ float32 metric_arg when metric_type > 1), uint64 count, count x float32 row-major.
    PARITY UNPINNED: Faiss is not installable in this image, so no real file was available to check this against."""
    import struct
    d, = struct.unpack("<i", f.read(4))
    ntotal, _d1, _d2 = struct.unpack("<qqq", f.read(24))
    _trained, = struct.unpack("<B", f.read(1))
    metric, = struct.unpack("<i", f.read(4))
    if metric > 1:
        f.read(4)
    count, = struct.unpack("<Q", f.read(8))
    if d <= 0 or ntotal < 0 or count != ntotal * d:
        raise ValueError(f"read_index: inconsistent Faiss flat header (d={d}, ntotal={ntotal}, payload={count} floats)")
    if metric not in (METRIC_INNER_PRODUCT, METRIC_L2):
        raise ValueError(f"read_index: Faiss metric_type {metric} is not served (inner product and L2 only)")
    if fourcc == b"IxFI":
        metric = METRIC_INNER_PRODUCT
    elif fourcc == b"IxF2":
        metric = METRIC_L2
    rows = np.fromfile(f, dtype="<f4", count=int(count)).reshape(int(ntotal), int(d))
    if rows.shape[0] != ntotal:
        raise ValueError("read_index: truncated Faiss flat file")
    return int(d), int(metric), rows


def write_index_faiss_flat(index, path: str) -> None:
    """The same rows in Faiss's native flat layout (see :func:`_read_faiss_flat`), for files that real Faiss should open."""
    import struct
    rows = index.reconstruct_n() if index.ntotal else np.empty((0, index.d), np.float32)
    with open(path, "wb") as f:
        f.write(b"IxFI" if index.metric_type == METRIC_INNER_PRODUCT else b"IxF2")
        f.write(struct.pack("<iqqqBi", int(index.d), int(rows.shape[0]), 1 << 20, 1 << 20, 1, int(index.metric_type)))
        f.write(struct.pack("<Q", int(rows.size)))
        np.ascontiguousarray(rows, dtype="<f4").tofile(f)


def read_index(path: str):
    """Reads files written by :func:`write_index`, ``.npy`` embedding caches (doc.npy / corpus_embeddings.npy,
    searcher.py:126-127,144) as inner-product indices, and native Faiss FLAT index files (serve/inference_service_text_retrieval.py:35
    reads whatever ``faiss.write_index`` produced; other Faiss index types need Faiss itself)."""
    with open(path, "rb") as f:
        head = f.read(8)
        if head == _MAGIC:
            d, metric, n = np.fromfile(f, dtype=np.int64, count=3)
            rows = np.fromfile(f, dtype=np.float32, count=int(n) * int(d)).reshape(int(n), int(d))
            idx = _LazyFlat(int(d), int(metric))
            if n:
                idx.add(rows)
            return idx
        if head[:4] in (b"IxFI", b"IxF2", b"IxFl"):
            f.seek(4)
            d, metric, rows = _read_faiss_flat(f, head[:4])
            idx = _LazyFlat(d, metric)
            if rows.shape[0]:
                idx.add(rows)
            return idx
    if head[:6] == b"\x93NUMPY":
        rows = np.load(path, mmap_mode="r")
        idx = _LazyFlat(int(rows.shape[1]), METRIC_INNER_PRODUCT)
        idx.add(np.ascontiguousarray(rows))
        return idx
    raise ValueError(f"read_index: {path} is neither a nexus_b200 flat index, a Faiss flat index, nor a .npy embedding cache "
                     "(other native Faiss index types are not readable without Faiss)")


def load_embedding_cache(path: str):
    """The embedding caches the reference writes, as ``(rows, ids)`` without touching the GPU:

    * ``.npy``  -- ``doc.npy`` / ``corpus_embeddings.npy`` (text_retrieval/searcher.py:126-127,144; memory-mapped) -> ``(rows, None)``
    * ``.pt``   -- ``item_vectors.pt`` = ``{'item_vectors': [N, D], 'item_ids': [N]}`` (training/.../trainer.py:38-39, read by
                   Nexus/inference/utils.py:9-12) -> ``(rows, item_ids)``
    * pickle    -- MMEB: an ``np.ndarray [Nq, D]`` (queries) or ``{candidate id: [D]}`` (eval_embedding.py:269-270,309-310,337-340)
                   -> ``(rows, None)`` / ``(rows, list of keys in dictionary order)``
    """
    with open(path, "rb") as f:
        head = f.read(8)
    if head[:6] == b"\x93NUMPY":
        return np.load(path, mmap_mode="r"), None
    if head[:2] == b"PK" or path.endswith(".pt"):
        import torch
        data = torch.load(path, map_location="cpu")                      # utils.py:9
        if not (isinstance(data, dict) and "item_vectors" in data):
            raise ValueError(f"{path}: a .pt cache must hold {{'item_vectors', 'item_ids'}} (got {type(data).__name__})")
        ids = data.get("item_ids")
        return data["item_vectors"].detach().cpu().numpy(), (ids.cpu().numpy() if ids is not None else None)   # :11-13
    import pickle
    with open(path, "rb") as f:
        obj = pickle.load(f)                                             # eval_embedding.py:338,340
    if isinstance(obj, dict):
        keys = list(obj.keys())
        if not keys:
            raise ValueError(f"{path}: empty candidate dictionary")
        return np.stack([np.asarray(obj[k]) for k in keys]), keys        # :353-354
    return np.asarray(obj), None


def index_from_cache(path: str, metric: int = METRIC_INNER_PRODUCT):
    """``(index, ids)``: a flat index over the rows of an embedding cache file (see :func:`load_embedding_cache`)."""
    rows, ids = load_embedding_cache(path)
    idx = _LazyFlat(int(rows.shape[1]), metric)
    idx.add(np.ascontiguousarray(rows))
    return idx, ids


def install_as_faiss(force: bool = False) -> types.ModuleType:
    """Register this module as ``faiss`` (and ``faiss.contrib.torch_utils``) in ``sys.modules`` so that
    the reference's ``import faiss`` picks up the B200 engine.  Does nothing if a real faiss is importable
    unless ``force``."""
    if not force:
        try:
            import faiss as real  # noqa: F401
            if not getattr(real, "__nexus_b200__", False):
                return real
        except ImportError:
            pass
    me = sys.modules[__name__]
    me.__nexus_b200__ = True
    contrib = types.ModuleType("faiss.contrib")
    torch_utils = types.ModuleType("faiss.contrib.torch_utils")
    contrib.torch_utils = torch_utils
    me.contrib = contrib
    sys.modules["faiss"] = me
    sys.modules["faiss.contrib"] = contrib
    sys.modules["faiss.contrib.torch_utils"] = torch_utils
    return me
