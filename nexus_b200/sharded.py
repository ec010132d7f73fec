"""Row-sharded multi-GPU flat search.

The reference delegates this to Faiss (``index_cpu_to_all_gpus`` with ``co.shard = True``,
Nexus/evaluation/text_retrieval/utils.py:178-183): every GPU holds N/G rows, every query goes to every
shard, per-shard top-k lists are merged on the host.  Here the merge stays on the device:

  * :class:`ShardedFlatIndex`      one process driving G devices (what ``index()`` builds when several GPUs
                                   are visible, like the reference): every shard writes its ``[Q,k]`` list as
                                   8-byte ``(score, shard-local row)`` entries (``nxf_search_packed``), the lists
                                   are copied peer-to-peer over NVLink to the home device (stream-ordered by
                                   events, no host synchronisation) and merged by ``nxf_merge_packed``.
  * :class:`DistributedFlatIndex`  one process per GPU (torchrun): the corpus rows never move after ``add``;
                                   the packed per-shard lists are exchanged with ONE NCCL all-gather of
                                   Q*k*8 bytes per rank into a preallocated buffer and merged by the same kernel
                                   on every rank, which adds each shard's row offset.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import numpy as np

from .flat import FlatIndex, _is_tensor, merge_packed, merge_topk, packed_padding


def shard_bounds(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous row ranges [lo, hi) of ``n`` rows over ``world`` shards; the first ``n % world``
    shards get one extra row.  Shard g owns global ids [lo_g, hi_g)."""
    base, rem = divmod(int(n), int(world))
    out, lo = [], 0
    for g in range(world):
        hi = lo + base + (1 if g < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


class ShardedFlatIndex:
    """One process, several devices (Faiss ``IndexShards`` shape)."""

    def __init__(self, d: int, metric: str = "ip", store_dtype: str = "bf16", devices: Optional[Sequence[int]] = None):
        import torch
        if devices is None:
            devices = list(range(torch.cuda.device_count()))
        if not devices:
            raise RuntimeError("ShardedFlatIndex: no CUDA devices")
        self.devices = [int(x) for x in devices]
        self.shards = [FlatIndex(d, metric, store_dtype, device=dev) for dev in self.devices]
        self._d = int(d)
        self._metric = metric
        self.nprobe = 1
        self.is_trained = True

    @property
    def d(self) -> int:
        return self._d

    @property
    def ntotal(self) -> int:
        return sum(s.ntotal for s in self.shards)

    @property
    def metric(self) -> str:
        return self._metric

    def train(self, x=None) -> None:
        return None

    def reset(self) -> None:
        for s in self.shards:
            s.reset()

    def add(self, x) -> None:
        """First add: rows split contiguously over the devices.  Later adds extend the last shard (ids stay
        insertion-ordered; the reference adds exactly once, utils.py:189)."""
        n = int(x.shape[0])
        if self.ntotal == 0:
            for (lo, hi), shard in zip(shard_bounds(n, len(self.shards)), self.shards):
                shard.set_id_offset(lo)
                if hi > lo:
                    part = x[lo:hi]
                    if _is_tensor(part) and part.is_cuda and part.device.index != shard.device:
                        part = part.to(f"cuda:{shard.device}")
                    shard.add(part)
        else:
            last = self.shards[-1]
            if last.ntotal == 0:
                last.set_id_offset(self.ntotal)
            part = x
            if _is_tensor(part) and part.is_cuda and part.device.index != last.device:
                part = part.to(f"cuda:{last.device}")
            last.add(part)

    def search(self, x, k: int):
        """Every shard searches on its own device (asynchronous, packed output); the lists reach the home device over NVLink
        behind events -- the host never waits for a shard -- and one merge kernel produces ``(D, I)`` with global ids."""
        import torch
        k = int(k)
        was_numpy = not _is_tensor(x)
        if was_numpy:
            a = np.asarray(x)
            xt = torch.from_numpy(np.ascontiguousarray(a if a.dtype in (np.float16, np.float32) else a.astype(np.float32)))
        else:
            xt = x
        if xt.dim() == 1:
            xt = xt.unsqueeze(0)
        was_cuda = (not was_numpy) and xt.is_cuda
        home = self.devices[0] if not was_cuda else xt.device.index
        nq = int(xt.shape[0])
        metric = "l2" if self._metric == "l2" else "ip"
        active = [s_ for s_ in self.shards if s_.ntotal > 0]
        if not active:
            fill = torch.finfo(torch.float32).max * (1 if metric == "l2" else -1)
            D = torch.full((nq, k), fill)
            I = torch.full((nq, k), -1, dtype=torch.int64)
            return (D.numpy(), I.numpy()) if was_numpy else (D.to(xt.device), I.to(xt.device))
        src_stream = torch.cuda.current_stream(xt.device) if was_cuda else None
        parts, events = [], []
        for shard in active:
            with torch.cuda.device(shard.device):
                st = torch.cuda.current_stream(shard.device)
                if src_stream is not None and shard.device != xt.device.index:
                    st.wait_stream(src_stream)                  # the queries were produced on the home device's stream
                q = xt.to(f"cuda:{shard.device}", non_blocking=True)
                P = shard.search_packed(q, k)
                ev = torch.cuda.Event()
                ev.record(st)
            parts.append(P)
            events.append(ev)
        with torch.cuda.device(home):
            hs = torch.cuda.current_stream(home)
            lists = torch.empty((len(active), nq, k), dtype=torch.int64, device=f"cuda:{home}")
            for g, (P, ev) in enumerate(zip(parts, events)):
                hs.wait_event(ev)
                lists[g].copy_(P, non_blocking=True)            # NVLink peer copy (a local copy for the home shard)
                P.record_stream(hs)
            D, I = merge_packed(lists, [s_.id_offset for s_ in active], k, metric)
        if was_numpy:
            return D.cpu().numpy(), I.cpu().numpy()
        if not was_cuda:
            return D.cpu(), I.cpu()
        return D, I


class DistributedFlatIndex:
    """One process per GPU over ``torch.distributed`` (NCCL).  Rank r holds shard r."""
    __nexus_b200_engine__ = True

    def __init__(self, d: int, metric: str = "ip", store_dtype: str = "bf16", device=None, group=None):
        import torch
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("DistributedFlatIndex needs an initialised torch.distributed process group")
        self.group = group
        self.rank = dist.get_rank(group)
        self.world = dist.get_world_size(group)
        if device is None:
            device = torch.cuda.current_device()
        self.local = FlatIndex(d, metric, store_dtype, device=device)
        self._metric = metric
        self._ntotal_global = 0
        self._offsets = [0] * self.world       # first global row of every rank's shard
        self._bufs = {}                        # (Q, k) -> preallocated exchange buffers

    @property
    def ntotal(self) -> int:
        return self._ntotal_global

    @property
    def d(self) -> int:
        return self.local.d

    def add_global(self, x) -> None:
        """Every rank passes the SAME [N, D] rows (or a lazily sliceable view of them, e.g. an np.memmap);
        rank r keeps rows shard_bounds(N, world)[r] and nothing else."""
        if self._ntotal_global != 0:
            raise RuntimeError("add_global: the sharded index is already populated (reset() first)")
        n = int(x.shape[0])
        lo, hi = shard_bounds(n, self.world)[self.rank]
        self.local.set_id_offset(lo)
        if hi > lo:
            self.local.add(x[lo:hi])
        self._ntotal_global = n
        self._offsets = [b[0] for b in shard_bounds(n, self.world)]

    def add_local(self, x_local, global_offset: int, n_global: int) -> None:
        """Rank-local rows that already are this rank's shard (e.g. straight from a data-parallel encoder:
        the embeddings are never all-gathered, unlike mmeb_v2/eval_embedding.py:127-131)."""
        import torch.distributed as dist
        self.local.set_id_offset(int(global_offset))
        self.local.add(x_local)
        self._ntotal_global = int(n_global)
        offs = [None] * self.world
        dist.all_gather_object(offs, int(global_offset), group=self.group)     # once per build: every rank's first global row
        self._offsets = [int(o) for o in offs]

    def reset(self) -> None:
        self.local.reset()
        self._ntotal_global = 0

    def _buffers(self, Q: int, k: int, device):
        import torch
        key = (Q, k)
        b = self._bufs.get(key)
        if b is None:
            if len(self._bufs) > 8:
                self._bufs.clear()
            b = (torch.empty((Q, k), dtype=torch.int64, device=device), torch.empty((self.world, Q, k), dtype=torch.int64, device=device))
            self._bufs[key] = b
        return b

    def search(self, q, k: int, gather=None, merge=None, exchange: str = "auto"):
        """``q``: the same [Q, D] CUDA tensor on every rank.  Returns the merged ``(D, I)`` on every rank.
        ``exchange``: ``"gather"`` = ONE all-gather of the per-shard ``[Q, k]`` lists as 8-byte ``(score, shard-local row)``
        entries into a preallocated buffer + the merge kernel on every rank, which adds the shards' row offsets (the north
        star's form); ``"partition"`` = rank r receives only the lists of its own block of queries (all-to-all), merges them and
        the final blocks are all-gathered -- 2/world of the bytes, for large ``Q*k``; ``"auto"`` picks by size.
        ``gather`` / ``merge`` are injection points for the CPU (gloo) tests of the exchange logic (unpacked lists)."""
        import torch
        import torch.distributed as dist
        k = int(k)
        metric = "l2" if self._metric == "l2" else "ip"
        if gather is None and merge is None and not (_is_tensor(q) and q.is_cuda):
            a = q if _is_tensor(q) else torch.from_numpy(np.ascontiguousarray(np.asarray(q)))
            q = (a if a.dtype in (torch.float16, torch.bfloat16, torch.float32) else a.float()).to(f"cuda:{self.local.device}")
        if gather is not None or merge is not None:
            D, I = self.local.search(q, k)
            if exchange == "partition":
                return exchange_partitioned(D, I, self.world, k, metric, self.group, merge or merge_topk)
            Ds, Is = (gather or all_gather_lists)(D, I, self.world, self.group)
            return (merge or merge_topk)(Ds, Is, k, metric)
        Q = int(q.shape[0])
        if exchange == "auto":
            exchange = "partition" if (self.world > 2 and Q * k * 8 * self.world >= PARTITION_MIN_BYTES) else "gather"
        if exchange == "partition":
            P = self.local.search_packed(q, k)
            return exchange_partitioned_packed(P, self._offsets, self.world, k, metric, self.group)
        mine, lists = self._buffers(Q, k, q.device)
        self.local.search_packed(q, k, out=mine)
        dist.all_gather_into_tensor(lists, mine, group=self.group)          # the one exchange step: Q*k*8 bytes per rank
        return merge_packed(lists, self._offsets, k, metric)


# all-gather volume (bytes received per rank) from which the query-partitioned exchange is used: below, one collective wins
PARTITION_MIN_BYTES = 256 << 20


def exchange_partitioned(D, I, world: int, k: int, metric: str = "ip", group=None, merge=None):
    """Query-partitioned exchange of per-shard top-k lists: ``[Q, k]`` on every rank -> merged ``[Q, k]`` on every rank.
    Queries are cut into ``world`` equal blocks (the last one padded with empty lists); one all-to-all hands rank r every shard's
    lists of block r, rank r merges them (``merge(Ds [world, Qb, k], Is, k, metric)``), one all-gather distributes the merged
    blocks.  Bytes received per rank: 2 * Q*k*12 * (world-1)/world instead of Q*k*12 * (world-1)."""
    import torch
    import torch.distributed as dist
    merge = merge or merge_topk
    Q = int(D.shape[0])
    Qb = (Q + world - 1) // world
    if Qb * world != Q:
        pad = Qb * world - Q
        fill = float(np.finfo(np.float32).max) if metric == "l2" else -float(np.finfo(np.float32).max)
        D = torch.cat([D, torch.full((pad, k), fill, dtype=D.dtype, device=D.device)], 0)
        I = torch.cat([I, torch.full((pad, k), -1, dtype=I.dtype, device=I.device)], 0)
    D, I = D.contiguous(), I.contiguous()
    Dr, Ir = torch.empty_like(D), torch.empty_like(I)
    dist.all_to_all_single(Dr, D, group=group)            # Dr[g*Qb:(g+1)*Qb] = shard g's lists for my query block
    dist.all_to_all_single(Ir, I, group=group)
    Dm, Im = merge(Dr.view(world, Qb, k), Ir.view(world, Qb, k), k, metric)
    Dm, Im = Dm.contiguous(), Im.contiguous()
    Do = torch.empty((world * Qb, k), dtype=Dm.dtype, device=Dm.device)
    Io = torch.empty((world * Qb, k), dtype=Im.dtype, device=Im.device)
    if hasattr(dist, "all_gather_into_tensor") and Dm.is_cuda:
        dist.all_gather_into_tensor(Do, Dm, group=group)
        dist.all_gather_into_tensor(Io, Im, group=group)
    else:
        dist.all_gather(list(Do.view(world, Qb, k).unbind(0)), Dm, group=group)
        dist.all_gather(list(Io.view(world, Qb, k).unbind(0)), Im, group=group)
    return Do[:Q], Io[:Q]


def exchange_partitioned_packed(P, offsets, world: int, k: int, metric: str = "ip", group=None):
    """:func:`exchange_partitioned` on packed lists (int64 ``[Q, k]`` of ``nxf_entry``): one all-to-all of 8-byte entries, the merge
    kernel adds the shard offsets, then the merged ``(D, I)`` blocks are all-gathered."""
    import torch
    import torch.distributed as dist
    Q = int(P.shape[0])
    Qb = (Q + world - 1) // world
    if Qb * world != Q:
        P = torch.cat([P, torch.full((Qb * world - Q, k), packed_padding(metric), dtype=P.dtype, device=P.device)], 0)
    P = P.contiguous()
    Pr = torch.empty_like(P)
    dist.all_to_all_single(Pr, P, group=group)           # Pr[g*Qb:(g+1)*Qb] = shard g's lists for my query block
    Dm, Im = merge_packed(Pr.view(world, Qb, k), offsets, k, metric)
    Do = torch.empty((world * Qb, k), dtype=Dm.dtype, device=Dm.device)
    Io = torch.empty((world * Qb, k), dtype=Im.dtype, device=Im.device)
    dist.all_gather_into_tensor(Do, Dm.contiguous(), group=group)
    dist.all_gather_into_tensor(Io, Im.contiguous(), group=group)
    return Do[:Q], Io[:Q]


def all_gather_lists(D, I, world: int, group=None):
    """ONE collective per tensor: [Q,k] -> [world, Q,k].  NCCL on GPUs; gloo in the CPU tests."""
    import torch
    import torch.distributed as dist
    Ds = torch.empty((world,) + tuple(D.shape), dtype=D.dtype, device=D.device)
    Is = torch.empty((world,) + tuple(I.shape), dtype=I.dtype, device=I.device)
    if hasattr(dist, "all_gather_into_tensor") and D.is_cuda:
        dist.all_gather_into_tensor(Ds, D.contiguous(), group=group)
        dist.all_gather_into_tensor(Is, I.contiguous(), group=group)
    else:
        dist.all_gather(list(Ds.unbind(0)), D.contiguous(), group=group)
        dist.all_gather(list(Is.unbind(0)), I.contiguous(), group=group)
    return Ds, Is
