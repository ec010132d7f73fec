"""An oracle-backed stand-in for libnexus_flat.so, for CPU tests of the host-side mirror ONLY (tests/test_reference_callers_cpu.py).

Same entry points, argument order and pointer conventions as include/nexus_flat.h, host pointers only; the arithmetic is the
oracle's restatement of the reference (NumpyIPIndex / blocked L2), so the real callers of the reference can be run end to end in
this GPU-less container through nexus_b200's Python layer -- patch.install(), retrieval.index/search, faiss_shim, flat.FlatIndex's
argument marshalling -- and compared with the unpatched reference.  It is test infrastructure: nothing under nexus_b200/ knows it.
"""
import collections
import ctypes as C

import numpy as np

import oracle

F32, F16, BF16 = 0, 1, 2


def _rows_from_ptr(ptr, n, ld, dim, dtype):
    if n == 0:
        return np.empty((0, dim), np.float32)
    if dtype == F32:
        a = np.ctypeslib.as_array((C.c_float * (n * ld)).from_address(ptr)).reshape(n, ld)[:, :dim]
        return np.array(a, dtype=np.float32)
    raw = np.ctypeslib.as_array((C.c_uint16 * (n * ld)).from_address(ptr)).reshape(n, ld)[:, :dim]
    if dtype == F16:
        return np.array(raw).view(np.float16).astype(np.float32)
    return (np.array(raw).astype(np.uint32) << 16).view(np.float32)


class FakeLib:
    def __init__(self):
        self.h = {}
        self.calls = collections.Counter()
        self._next = 1
        self._err = b""

    # ---- library ----
    def nxf_version(self):
        return b"fake (oracle-backed, CPU tests only)"

    def nxf_last_error(self):
        return self._err

    def nxf_device_count(self):
        return 1

    # ---- handle ----
    def nxf_create(self, dim, metric, store, device, out):
        self.calls["create"] += 1
        hid = self._next
        self._next += 1
        self.h[hid] = {"dim": dim, "metric": metric, "store": store, "device": device, "rows": np.empty((0, dim), np.float32), "off": 0}
        out._obj.value = hid
        return 0

    def _get(self, h):
        return self.h[h.value if hasattr(h, "value") else h]

    def nxf_destroy(self, h):
        self.h.pop(h.value if hasattr(h, "value") else h, None)

    def nxf_reset(self, h):
        x = self._get(h)
        x["rows"] = np.empty((0, x["dim"]), np.float32)
        return 0

    def nxf_reserve(self, h, n):
        return 0

    def nxf_ntotal(self, h):
        return self._get(h)["rows"].shape[0]

    def nxf_dim(self, h):
        return self._get(h)["dim"]

    def nxf_metric(self, h):
        return self._get(h)["metric"]

    def nxf_store_dtype(self, h):
        return self._get(h)["store"]

    def nxf_device(self, h):
        return self._get(h)["device"]

    def nxf_set_id_offset(self, h, off):
        self._get(h)["off"] = off
        return 0

    # ---- data ----
    def nxf_add(self, h, ptr, n, dtype, ld, on_device, stream):
        self.calls["add"] += 1
        x = self._get(h)
        assert not on_device, "the fake library takes host pointers only"
        rows = _rows_from_ptr(ptr, n, ld or x["dim"], x["dim"], dtype)
        if x["store"] == BF16:
            rows = oracle.round_to_bf16(rows)
        elif x["store"] == F16:
            rows = oracle.round_to_f16(rows)
        x["rows"] = np.concatenate([x["rows"], rows], axis=0)
        return 0

    def nxf_reconstruct(self, h, first, n, dst, on_device, stream):
        x = self._get(h)
        out = np.ctypeslib.as_array((C.c_float * (n * x["dim"])).from_address(dst)).reshape(n, x["dim"])
        out[:] = x["rows"][first:first + n]
        return 0

    def nxf_search(self, h, qptr, nq, qdtype, qld, q_on_device, k, dptr, iptr, out_on_device, flags, stream):
        self.calls["search"] += 1
        self.calls["search_rows"] += nq
        x = self._get(h)
        assert not q_on_device and not out_on_device, "the fake library takes host pointers only"
        if nq == 0:
            return 0
        q = _rows_from_ptr(qptr, nq, qld or x["dim"], x["dim"], qdtype)
        if x["metric"] == 0:
            s, i = oracle.numpy_ip_search(x["rows"], q, k)          # the reference's own arithmetic and order (utils.py:16-31)
            s, i = oracle.faiss_pad(s, i, k)
        else:
            s, i = oracle.blocked_l2_search(x["rows"], q, k)
        i = np.where(i >= 0, i + x["off"], -1)
        np.ctypeslib.as_array((C.c_float * (nq * k)).from_address(dptr)).reshape(nq, k)[:] = s
        np.ctypeslib.as_array((C.c_int64 * (nq * k)).from_address(iptr)).reshape(nq, k)[:] = i
        return 0
