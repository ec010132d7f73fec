"""FlatIndex: the Python face of one device-resident exact flat index (one GPU).

Mirrors the object surface the reference touches on a Faiss flat index:
``.add(x)``, ``.train(x)`` (no-op), ``.search(x, k) -> (D float32 [n,k], I int64 [n,k])``, ``.ntotal``, ``.d``,
a settable-and-ignored ``.nprobe`` (Nexus/inference/embedder/recommendation/base.py:117), ``.reset()``,
``.reconstruct_n()``.  Call sites: Nexus/evaluation/text_retrieval/utils.py:188-189,224;
Nexus/training/embedder/recommendation/modeling.py:195,197; serve/inference_service_text_retrieval.py:70.

numpy in -> numpy out (host buffers, synchronous); torch CUDA tensor in -> torch CUDA tensors out on the
same device, asynchronous on the current stream (the ``faiss.contrib.torch_utils`` behaviour that
modeling.py:9 relies on); torch CPU tensor in -> torch CPU tensors out.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Tuple, Union

import numpy as np

from . import _lib as L

_NP_DTYPES = {np.dtype(np.float32): L.F32, np.dtype(np.float16): L.F16}
_STORE_NAMES = {"f32": L.F32, "fp32": L.F32, "float32": L.F32, "f16": L.F16, "fp16": L.F16, "float16": L.F16,
                "half": L.F16, "bf16": L.BF16, "bfloat16": L.BF16}
_METRICS = {"ip": L.METRIC_IP, "inner_product": L.METRIC_IP, "cosine": L.METRIC_IP, "l2": L.METRIC_L2}


def _torch():
    import torch
    return torch


def _is_tensor(x) -> bool:
    return type(x).__module__.startswith("torch") and hasattr(x, "data_ptr")


def parse_device(device) -> int:
    """None / int / 'cuda' / 'cuda:3' / torch.device -> CUDA ordinal."""
    if device is None:
        try:
            torch = _torch()
            return torch.cuda.current_device() if torch.cuda.is_available() else 0
        except Exception:
            return 0
    if isinstance(device, int):
        return device
    s = str(device)
    if s in ("cuda", "gpu"):
        return parse_device(None)
    if s.startswith("cuda:"):
        return int(s.split(":", 1)[1])
    if s.isdigit():
        return int(s)
    raise ValueError(f"nexus_b200 indices live on a CUDA device; got device={device!r} (there is no CPU path)")


def parse_store_dtype(dt) -> int:
    if isinstance(dt, int):
        return dt
    s = str(dt).replace("torch.", "").replace("numpy.", "").lower()
    if s in _STORE_NAMES:
        return _STORE_NAMES[s]
    raise ValueError(f"unknown storage dtype {dt!r} (use 'bf16', 'f16' or 'f32')")


class _Buf:
    """A (pointer, dtype code, row stride, on_device, keepalive) view of a 2-D row-major array."""
    __slots__ = ("ptr", "dtype", "ld", "on_device", "n", "d", "keep", "device", "kind")


def _as_rows(x, want_dim: Optional[int], what: str) -> _Buf:
    b = _Buf()
    if _is_tensor(x):
        torch = _torch()
        if x.dim() == 1:
            x = x.unsqueeze(0)
        if x.dim() != 2:
            raise ValueError(f"{what}: expected a 2-D tensor, got shape {tuple(x.shape)}")
        codes = {torch.float32: L.F32, torch.float16: L.F16, torch.bfloat16: L.BF16}
        if x.dtype not in codes:
            x = x.float()
        if x.stride(1) != 1 or (x.shape[0] > 1 and x.stride(0) < x.shape[1]):
            x = x.contiguous()
        elif not x.is_cuda and x.shape[0] > 1 and x.stride(0) != x.shape[1]:
            # a strided HOST view (big[:, off:off+d]): the host paths copy n*ld contiguous elements from data_ptr, which would
            # run past the end of the storage on the last row -- densify (device views are read element-wise and may stay)
            x = x.contiguous()
        b.ptr, b.dtype, b.ld = x.data_ptr(), codes[x.dtype], (x.stride(0) if x.shape[0] > 1 else x.shape[1])
        b.on_device = 1 if x.is_cuda else 0
        b.device = x.device.index if x.is_cuda else None
        b.n, b.d, b.keep, b.kind = x.shape[0], x.shape[1], x, "torch"
    else:
        a = np.asarray(x)
        if a.ndim == 1:
            a = a[None, :]
        if a.ndim != 2:
            raise ValueError(f"{what}: expected a 2-D array, got shape {a.shape}")
        if a.dtype not in _NP_DTYPES:
            a = a.astype(np.float32)           # the reference casts everything to fp32 (utils.py:169,224)
        a = np.ascontiguousarray(a)
        b.ptr, b.dtype, b.ld, b.on_device, b.device = a.ctypes.data, _NP_DTYPES[a.dtype], a.shape[1], 0, None
        b.n, b.d, b.keep, b.kind = a.shape[0], a.shape[1], a, "numpy"
    if want_dim is not None and b.d != want_dim:
        raise ValueError(f"{what}: rows have {b.d} columns, index dimension is {want_dim}")
    return b


def _stream_ptr(device: int) -> int:
    try:
        torch = _torch()
        if torch.cuda.is_available():
            return int(torch.cuda.current_stream(device).cuda_stream)
    except Exception:
        pass
    return 0


class FlatIndex:
    """Exact inner-product / L2 index resident on one B200.

    Args:
        d: embedding dimension.
        metric: 'ip' (also what cosine reduces to on normalised embeddings) or 'l2'.
        store_dtype: 'bf16' (default for the tensor-core path), 'f16' or 'f32'.
        device: CUDA ordinal / 'cuda:N' / None (current device).
    """

    def __init__(self, d: int, metric: Union[str, int] = "ip", store_dtype: Union[str, int] = "bf16", device=None):
        lib = L.load()
        self._lib = lib
        self._h = C.c_void_p()
        m = metric if isinstance(metric, int) else _METRICS[str(metric).lower()]
        self._device = parse_device(device)
        L.check(lib.nxf_create(int(d), int(m), parse_store_dtype(store_dtype), self._device, C.byref(self._h)))
        self.nprobe = 1          # settable and ignored: an exact scan probes everything
        self.is_trained = True
        self.search_flags = L.SEARCH_DEFAULT
        self._id_offset = 0

    # ---- Faiss-shaped attributes ----
    @property
    def d(self) -> int:
        return self._lib.nxf_dim(self._h)

    @property
    def ntotal(self) -> int:
        return int(self._lib.nxf_ntotal(self._h))

    @property
    def metric(self) -> str:
        return "l2" if self._lib.nxf_metric(self._h) == L.METRIC_L2 else "ip"

    @property
    def metric_type(self) -> int:
        return self._lib.nxf_metric(self._h)

    @property
    def store_dtype(self) -> str:
        return {L.F32: "f32", L.F16: "f16", L.BF16: "bf16"}[self._lib.nxf_store_dtype(self._h)]

    @property
    def device(self) -> int:
        return self._device

    def __len__(self) -> int:
        return self.ntotal

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            try:
                self._lib.nxf_destroy(h)
            except Exception:
                pass

    # ---- build ----
    def train(self, x=None) -> None:
        """No-op: a flat index has nothing to train (utils.py:188)."""
        return None

    def reserve(self, n_total: int) -> None:
        L.check(self._lib.nxf_reserve(self._h, int(n_total)))

    def set_id_offset(self, offset: int) -> None:
        L.check(self._lib.nxf_set_id_offset(self._h, int(offset)))
        self._id_offset = int(offset)

    @property
    def id_offset(self) -> int:
        return self._id_offset

    def add(self, x) -> None:
        """Append rows (utils.py:189, modeling.py:195).  The index keeps its own converted copy."""
        b = _as_rows(x, self.d, "add")
        if b.on_device and b.device != self._device:
            raise ValueError(f"add: rows live on cuda:{b.device}, index on cuda:{self._device}")
        L.check(self._lib.nxf_add(self._h, b.ptr, b.n, b.dtype, b.ld, b.on_device, _stream_ptr(self._device)))

    def reset(self) -> None:
        L.check(self._lib.nxf_reset(self._h))

    def reconstruct_n(self, first: int = 0, n: Optional[int] = None) -> np.ndarray:
        n = self.ntotal - first if n is None else n
        out = np.empty((n, self.d), dtype=np.float32)
        L.check(self._lib.nxf_reconstruct(self._h, int(first), int(n), out.ctypes.data, 0, _stream_ptr(self._device)))
        return out

    def reconstruct_into(self, out, first: int = 0) -> None:
        """Rows ``[first, first + out.shape[0])`` as fp32 into a CUDA tensor on the index's device (no host round trip)."""
        torch = _torch()
        if not (out.is_cuda and out.device.index == self._device and out.dtype == torch.float32 and out.is_contiguous()
                and out.dim() == 2 and out.shape[1] == self.d):
            raise ValueError(f"reconstruct_into: need a contiguous float32 CUDA tensor [n, {self.d}] on cuda:{self._device}")
        L.check(self._lib.nxf_reconstruct(self._h, int(first), int(out.shape[0]), out.data_ptr(), 1, _stream_ptr(self._device)))

    def reconstruct(self, i: int) -> np.ndarray:
        return self.reconstruct_n(i, 1)[0]

    # ---- search ----
    def search(self, x, k: int, flags: Optional[int] = None, out=None) -> Tuple:
        """``(D float32 [n,k], I int64 [n,k])``, rows best-first, ``-1`` padding when k > ntotal."""
        b = _as_rows(x, self.d, "search")
        k = int(k)
        flags = self.search_flags if flags is None else flags
        stream = _stream_ptr(self._device)
        if b.kind == "torch":
            torch = _torch()
            if b.on_device:
                if b.device != self._device:
                    raise ValueError(f"search: queries live on cuda:{b.device}, index on cuda:{self._device}")
                D = torch.empty((b.n, k), dtype=torch.float32, device=b.keep.device)
                I = torch.empty((b.n, k), dtype=torch.int64, device=b.keep.device)
                L.check(self._lib.nxf_search(self._h, b.ptr, b.n, b.dtype, b.ld, 1, k, D.data_ptr(), I.data_ptr(), 1, flags, stream))
                return D, I
            if out is not None:       # caller-provided (e.g. pinned) CPU result tensors
                D, I = out
                assert tuple(D.shape) == (b.n, k) and tuple(I.shape) == (b.n, k) and D.is_contiguous() and I.is_contiguous()
                assert D.dtype == torch.float32 and I.dtype == torch.int64 and not D.is_cuda and not I.is_cuda
            else:
                D = torch.empty((b.n, k), dtype=torch.float32)
                I = torch.empty((b.n, k), dtype=torch.int64)
            L.check(self._lib.nxf_search(self._h, b.ptr, b.n, b.dtype, b.ld, 0, k, D.data_ptr(), I.data_ptr(), 0, flags, stream))
            return D, I
        D = np.empty((b.n, k), dtype=np.float32)
        I = np.empty((b.n, k), dtype=np.int64)
        L.check(self._lib.nxf_search(self._h, b.ptr, b.n, b.dtype, b.ld, 0, k, D.ctypes.data, I.ctypes.data, 0, flags, stream))
        return D, I

    def search_packed(self, x, k: int, flags: Optional[int] = None, out=None):
        """The shard's contribution to a row-sharded search: ``x`` a CUDA tensor on this device -> int64 CUDA tensor ``[n, k]``
        whose 8-byte elements are ``(float32 score, int32 shard-local row)`` pairs (``nxf_entry``; row -1 = padding), rows
        best-first.  One buffer, so the per-shard lists travel in ONE collective; :func:`merge_packed` turns the gathered lists
        into ``(D, I)`` with global ids.  Asynchronous on the current stream."""
        torch = _torch()
        b = _as_rows(x, self.d, "search_packed")
        if b.kind != "torch" or not b.on_device or b.device != self._device:
            raise ValueError(f"search_packed: queries must be a CUDA tensor on cuda:{self._device}")
        k = int(k)
        flags = self.search_flags if flags is None else flags
        if out is None:
            out = torch.empty((b.n, k), dtype=torch.int64, device=b.keep.device)
        elif tuple(out.shape) != (b.n, k) or out.dtype != torch.int64 or not out.is_contiguous() or out.device != b.keep.device:
            raise ValueError("search_packed: out must be a contiguous int64 CUDA tensor [n, k] on the index's device")
        L.check(self._lib.nxf_search_packed(self._h, b.ptr, b.n, b.dtype, b.ld, k, out.data_ptr(), flags, _stream_ptr(self._device)))
        return out

    def search_subsets(self, x, offsets, rows, k: int) -> Tuple:
        """Per-query candidate lists: query ``i`` ranks the index rows ``rows[offsets[i]:offsets[i+1]]`` (CSR) and gets the
        best ``k`` POSITIONS inside its own list -> ``(D float32 [n,k], P int64 [n,k])``, best first, ties by ascending position,
        ``-1`` padding beyond the list length.  The device form of the local-ranking loop at
        Nexus/evaluation/mmeb_v2/eval_embedding.py:384-404.  ``offsets``/``rows`` int64, both numpy/CPU or both CUDA tensors."""
        b = _as_rows(x, self.d, "search_subsets")
        k = int(k)
        stream = _stream_ptr(self._device)
        lists_on_device = _is_tensor(offsets) and offsets.is_cuda
        if lists_on_device != (_is_tensor(rows) and rows.is_cuda):
            raise ValueError("search_subsets: offsets and rows must both be CUDA tensors or both host arrays")
        if lists_on_device:
            torch = _torch()
            if offsets.device.index != self._device or rows.device.index != self._device:
                raise ValueError(f"search_subsets: the lists must live on cuda:{self._device}")
            off = offsets.to(torch.int64).contiguous()
            rws = rows.to(torch.int64).contiguous()
            if off.dim() != 1 or off.numel() != b.n + 1:
                raise ValueError(f"search_subsets: offsets has {off.numel()} entries for {b.n} queries (want nq + 1)")
            total, off_ptr, rows_ptr = int(rws.numel()), off.data_ptr(), rws.data_ptr()
        else:
            off = np.ascontiguousarray(np.asarray(offsets.cpu() if _is_tensor(offsets) else offsets, dtype=np.int64))
            rws = np.ascontiguousarray(np.asarray(rows.cpu() if _is_tensor(rows) else rows, dtype=np.int64))
            if off.ndim != 1 or off.size != b.n + 1:
                raise ValueError(f"search_subsets: offsets has {off.size} entries for {b.n} queries (want nq + 1)")
            total, off_ptr, rows_ptr = int(rws.size), off.ctypes.data, rws.ctypes.data
        call = self._lib.nxf_search_subsets
        if b.kind == "torch":
            torch = _torch()
            if b.on_device and b.device != self._device:
                raise ValueError(f"search_subsets: queries live on cuda:{b.device}, index on cuda:{self._device}")
            dev = b.keep.device if b.on_device else "cpu"
            D = torch.empty((b.n, k), dtype=torch.float32, device=dev)
            P = torch.empty((b.n, k), dtype=torch.int64, device=dev)
            L.check(call(self._h, b.ptr, b.n, b.dtype, b.ld, b.on_device, off_ptr, rows_ptr, total, int(lists_on_device), k,
                         D.data_ptr(), P.data_ptr(), b.on_device, stream))
            return D, P
        D = np.empty((b.n, k), dtype=np.float32)
        P = np.empty((b.n, k), dtype=np.int64)
        L.check(call(self._h, b.ptr, b.n, b.dtype, b.ld, 0, off_ptr, rows_ptr, total, int(lists_on_device), k,
                     D.ctypes.data, P.ctypes.data, 0, stream))
        return D, P

    def set_profiling(self, enable: bool = True) -> None:
        """Time every main search kernel with CUDA events on its stream (read back through stats())."""
        L.check(self._lib.nxf_set_profiling(self._h, 1 if enable else 0))

    def stats(self) -> dict:
        st = L.Stats()
        L.check(self._lib.nxf_get_stats(self._h, C.byref(st)))
        out = {name: getattr(st, name) for name, _ in L.Stats._fields_}
        out["last_path_name"] = L.PATH_NAMES.get(st.last_path, "?")
        return out


def merge_topk(scores, ids, k_out: Optional[int] = None, metric: str = "ip"):
    """Merge per-shard sorted lists on the device: ``scores`` / ``ids`` are CUDA tensors
    ``[n_lists, nq, k]`` (float32 / int64 global ids, -1 = padding) -> ``([nq, k_out], [nq, k_out])``.
    The device-side replacement of Faiss IndexShards' host merge (utils.py:178-183)."""
    torch = _torch()
    if not (scores.is_cuda and ids.is_cuda):
        raise ValueError("merge_topk works on CUDA tensors only (there is no CPU path)")
    scores = scores.contiguous().float()
    ids = ids.contiguous().long()
    n_lists, nq, k = scores.shape
    k_out = k if k_out is None else int(k_out)
    D = torch.empty((nq, k_out), dtype=torch.float32, device=scores.device)
    I = torch.empty((nq, k_out), dtype=torch.int64, device=scores.device)
    dev = scores.device.index
    L.check(L.load().nxf_merge_topk(scores.data_ptr(), ids.data_ptr(), n_lists, nq, k, k_out, _METRICS[metric],
                                    D.data_ptr(), I.data_ptr(), dev, _stream_ptr(dev)))
    return D, I


def packed_padding(metric: str = "ip") -> int:
    """The int64 value of a padding ``nxf_entry`` (row -1, score -FLT_MAX for inner product / +FLT_MAX for L2)."""
    fill = np.finfo(np.float32).max if metric == "l2" else -np.finfo(np.float32).max
    return int(np.array([(fill, -1)], dtype=[("score", "<f4"), ("row", "<i4")]).view(np.int64)[0])


def merge_packed(lists, offsets, k_out: Optional[int] = None, metric: str = "ip", out=None):
    """Merge gathered per-shard packed lists on the device: ``lists`` int64 CUDA tensor ``[n_lists, nq, k]`` of ``nxf_entry``
    pairs (see :meth:`FlatIndex.search_packed`), ``offsets`` the first global row of every shard ->
    ``(D float32 [nq, k_out], I int64 [nq, k_out])`` with global ids.  K3 reading 8-byte entries: the device-side replacement of
    Faiss IndexShards' host merge (utils.py:178-183)."""
    torch = _torch()
    if not lists.is_cuda or lists.dtype != torch.int64 or lists.dim() != 3:
        raise ValueError("merge_packed: lists must be an int64 CUDA tensor [n_lists, nq, k] (there is no CPU path)")
    lists = lists.contiguous()
    n_lists, nq, k = lists.shape
    if len(offsets) != n_lists:
        raise ValueError(f"merge_packed: {len(offsets)} offsets for {n_lists} lists")
    k_out = k if k_out is None else int(k_out)
    if out is None:
        D = torch.empty((nq, k_out), dtype=torch.float32, device=lists.device)
        I = torch.empty((nq, k_out), dtype=torch.int64, device=lists.device)
    else:
        D, I = out
    offs = (C.c_int64 * n_lists)(*[int(o) for o in offsets])
    dev = lists.device.index
    L.check(L.load().nxf_merge_packed(lists.data_ptr(), offs, n_lists, nq, k, k_out, _METRICS[metric], D.data_ptr(), I.data_ptr(),
                                      dev, _stream_ptr(dev)))
    return D, I


def positive_rank(pos_scores, topk_scores, return_pred: bool = True):
    """Position of the positive among the top-k, on the device: the fused form of ``torch.cat([pos, topk])`` + stable
    descending ``torch.sort`` + ``indice[:, :k] == 0`` in ``BaseRetriever.eval_step``
    (Nexus/training/embedder/recommendation/modeling.py:199-205).

    ``pos_scores`` [B] or [B, 1] and ``topk_scores`` [B, k] CUDA tensors -> ``rank`` int32 [B] in [0, k] (k = the positive
    is behind all k) and, with ``return_pred``, ``pred`` bool [B, k] exactly as the reference's ``pred``."""
    torch = _torch()
    if not (pos_scores.is_cuda and topk_scores.is_cuda):
        raise ValueError("positive_rank works on CUDA tensors only (there is no CPU path)")
    topk = topk_scores.contiguous().float()
    pos = pos_scores.reshape(-1).contiguous().float()
    B, k = topk.shape
    if pos.shape[0] != B:
        raise ValueError(f"pos_scores has {pos.shape[0]} rows, topk_scores {B}")
    rank = torch.empty((B,), dtype=torch.int32, device=topk.device)
    pred = torch.empty((B, k), dtype=torch.uint8, device=topk.device) if return_pred else None
    dev = topk.device.index
    L.check(L.load().nxf_positive_rank(pos.data_ptr(), topk.data_ptr(), B, k, rank.data_ptr(),
                                       pred.data_ptr() if return_pred else 0, dev, _stream_ptr(dev)))
    return (rank, pred.bool()) if return_pred else rank
