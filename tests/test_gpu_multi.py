"""Row a7 on real devices: the two row-sharded index classes (nexus_b200/sharded.py) with >= 2 GPUs.

  * ShardedFlatIndex      one process, G devices -- what ``index()`` builds when several GPUs are visible, the shape of
                          ``faiss.index_cpu_to_all_gpus(index, co)`` with ``co.shard = True`` (utils.py:178-183);
  * DistributedFlatIndex  one process per GPU (spawned here, NCCL): packed per-shard lists, ONE all-gather, merge kernel.

Skipped below 2 visible devices (run with ``gpurun --gpus 2``; the log of that run is kept under profiles/)."""
import os
import socket

import numpy as np
import pytest

import oracle
from _common import gen_inputs

pytestmark = pytest.mark.gpu


def _ngpu():
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs >= 2 CUDA devices")


@needs2
@pytest.mark.parametrize("store,k,n,d,q", [("bf16", 100, 200001, 128, 333), ("f32", 10, 50000, 64, 70), ("f16", 500, 300000, 64, 40), ("bf16", 20, 7, 32, 5)])
def test_sharded_flat_index_numpy_and_cuda(store, k, n, d, q):
    import torch
    from nexus_b200 import ShardedFlatIndex
    G = torch.cuda.device_count()
    c, x = gen_inputs(n, d, q, seed=60 + d)
    idx = ShardedFlatIndex(d, "ip", store)
    idx.add(c)
    assert idx.ntotal == n and len(idx.shards) == G
    assert [s.ntotal for s in idx.shards] == [hi - lo for lo, hi in __import__("nexus_b200").shard_bounds(n, G)]
    cs = {"bf16": oracle.round_to_bf16, "f16": oracle.round_to_f16, "f32": lambda a: a}[store](c)
    ref_s, ref_i = oracle.blocked_ip_search(cs, x, k)
    D, I = idx.search(x, k)                                             # numpy in -> numpy out
    assert isinstance(D, np.ndarray) and D.shape == (q, k) and I.dtype == np.int64
    rep = oracle.check_topk(ref_s, ref_i, D, I, cs, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]
    for home in (0, G - 1):                                             # CUDA in -> CUDA out on the queries' device, no host sync inside
        xt = torch.from_numpy(x).to(f"cuda:{home}")
        Dt, It = idx.search(xt, k)
        assert Dt.device.index == home and It.device.index == home
        rep = oracle.check_topk(ref_s, ref_i, Dt.cpu().numpy(), It.cpu().numpy(), cs, x, rtol=1e-5)
        assert rep.ok, (home, rep.messages[:5])


@needs2
def test_index_builds_a_sharded_index_when_several_gpus_are_visible():
    """retrieval.index(device=None) follows utils.py:178-183: all visible GPUs, rows sharded; search() in one call."""
    import nexus_b200
    c, x = gen_inputs(120000, 96, 150, seed=71)
    idx = nexus_b200.index(corpus_embeddings=c)
    assert isinstance(idx, nexus_b200.ShardedFlatIndex)
    s, i = nexus_b200.search(idx, k=100, query_embeddings=x)
    ref_s, ref_i = oracle.blocked_ip_search(c, x, 100)
    rep = oracle.check_topk(ref_s, ref_i, s, i, c, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]
    idx.add(c[:1000] * 0.5)                                             # a later add extends the last shard, ids stay insertion-ordered
    assert idx.ntotal == 121000
    s2, i2 = nexus_b200.search(idx, k=100, query_embeddings=x)
    c2 = np.concatenate([c, c[:1000] * 0.5])
    ref_s, ref_i = oracle.blocked_ip_search(c2, x, 100)
    rep = oracle.check_topk(ref_s, ref_i, s2, i2, c2, x, rtol=1e-5)
    assert rep.ok, rep.messages[:5]


@needs2
def test_sharded_l2_metric():
    from nexus_b200 import ShardedFlatIndex
    c, x = gen_inputs(30000, 64, 50, seed=81, normalize=False)
    idx = ShardedFlatIndex(64, "l2", "f32")
    idx.add(c)
    D, I = idx.search(x, 10)
    ref_s, ref_i = oracle.blocked_l2_search(c, x, 10)
    rep = oracle.check_topk(ref_s, ref_i, D, I, None, None, rtol=1e-5, atol=1e-4, tie_tol=1e-3, descending=False)
    assert rep.ok, rep.messages[:5]


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _dist_worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    from nexus_b200 import DistributedFlatIndex, shard_bounds
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        c, x = gen_inputs(150001, 128, 777, seed=91)
        cs = oracle.round_to_bf16(c)
        idx = DistributedFlatIndex(128, "ip", "bf16")
        idx.add_global(c)                                                   # every rank passes the same rows, keeps its shard
        lo, hi = shard_bounds(c.shape[0], world)[rank]
        assert idx.local.ntotal == hi - lo and idx.ntotal == c.shape[0]
        q = torch.from_numpy(x).cuda()
        res = {}
        for mode in ("gather", "partition"):
            for k in (100, 10):
                D, I = idx.search(q, k, exchange=mode)
                D2, I2 = idx.search(q, k, exchange=mode)                    # buffers are reused: a second call must not corrupt the first result
                torch.cuda.synchronize()
                assert torch.equal(I, I2) and torch.equal(D, D2)
                res[(mode, k)] = (D.cpu().numpy(), I.cpu().numpy())
        # rank-local rows straight from a (pretend) data-parallel encoder: add_local
        idx2 = DistributedFlatIndex(128, "ip", "bf16")
        idx2.add_local(torch.from_numpy(c[lo:hi]).cuda(), lo, c.shape[0])
        D3, I3 = idx2.search(q, 100)
        torch.cuda.synchronize()
        res[("local", 100)] = (D3.cpu().numpy(), I3.cpu().numpy())
        if rank == 0:
            for (mode, k), (D, I) in res.items():
                ref_s, ref_i = oracle.blocked_ip_search(cs, x, k)
                rep = oracle.check_topk(ref_s, ref_i, D, I, cs, x, rtol=1e-5)
                assert rep.ok, (mode, k, rep.messages[:5])
        np.save(os.path.join(out_dir, f"ids_{rank}.npy"), res[("gather", 100)][1])
    finally:
        dist.destroy_process_group()


@needs2
def test_distributed_flat_index_nccl_two_ranks(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_dist_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    np.testing.assert_array_equal(np.load(tmp_path / "ids_0.npy"), np.load(tmp_path / "ids_1.npy"))


def test_packed_search_and_merge_on_logical_shards():
    """nxf_search_packed / nxf_merge_packed on ONE device (logical shards): equal to the unsharded search, for both metrics, with
    padding (a shard smaller than k) and an empty shard."""
    import torch
    from nexus_b200 import FlatIndex, merge_packed, shard_bounds
    from nexus_b200.flat import packed_padding
    for metric, store in (("ip", "bf16"), ("l2", "f32"), ("ip", "f32")):
        c, x = gen_inputs(20011, 64, 130, seed=17, normalize=(metric == "ip"))
        cs = oracle.round_to_bf16(c) if store == "bf16" else c
        xt = torch.from_numpy(x).cuda()
        k = 60
        bounds = shard_bounds(c.shape[0] - 25, 3) + [(c.shape[0] - 25, c.shape[0])]          # the last shard holds 25 < k rows
        lists, offs = [], []
        for lo, hi in bounds:
            s = FlatIndex(64, metric, store)
            s.set_id_offset(lo)
            s.add(c[lo:hi])
            lists.append(s.search_packed(xt, k))
            offs.append(lo)
        empty = torch.full((130, k), packed_padding(metric), dtype=torch.int64, device="cuda")   # what an empty shard contributes
        lists.append(empty)
        offs.append(0)
        D, I = merge_packed(torch.stack(lists), offs, k, metric)
        if metric == "ip":
            ref_s, ref_i = oracle.blocked_ip_search(cs, x, k)
            rep = oracle.check_topk(ref_s, ref_i, D.cpu().numpy(), I.cpu().numpy(), cs, x, rtol=1e-5)
        else:
            ref_s, ref_i = oracle.blocked_l2_search(cs, x, k)
            rep = oracle.check_topk(ref_s, ref_i, D.cpu().numpy(), I.cpu().numpy(), None, None, rtol=1e-5, atol=1e-4, tie_tol=1e-3, descending=False)
        assert rep.ok, (metric, store, rep.messages[:5])
        last = lists[3].cpu().numpy().view([("score", "<f4"), ("row", "<i4")])
        assert (last["row"][:, 25:] == -1).all() and (last["row"][:, :25] >= 0).all() and (last["row"][:, :25] < 25).all()   # shard-local rows


@pytest.mark.parametrize("G,k,k_out,metric", [(2, 100, 100, "ip"), (8, 500, 500, "ip"), (9, 500, 100, "ip"), (8, 100, 37, "l2"), (3, 7, 7, "ip"), (64, 60, 60, "ip")])
def test_merge_kernels_on_sorted_lists_with_ties_and_padding(G, k, k_out, metric):
    """K3 in both forms (rank merge of sorted lists up to 4 096 entries per query, select + sort beyond) against a numpy merge:
    equal scores across lists (order: ascending global id), lists shorter than k (-1 padding), fewer valid entries than k_out."""
    import torch
    from nexus_b200 import merge_topk
    rng = np.random.default_rng(G * 1000 + k)
    nq = 50
    pool = np.round(rng.standard_normal(400), 2).astype(np.float32)           # few distinct values: many exact ties
    S = np.empty((G, nq, k), np.float32)
    I = np.empty((G, nq, k), np.int64)
    sign = -1.0 if metric == "l2" else 1.0
    for g in range(G):
        for q in range(nq):
            n_valid = k if (g + q) % 5 else int(rng.integers(0, k + 1))
            sc = rng.choice(pool, size=k).astype(np.float32)
            if metric == "l2":
                sc = np.abs(sc)
            ids = (rng.choice(10_000, size=k, replace=False) * G + g).astype(np.int64)   # disjoint id sets per list, like row shards
            order = np.lexsort((ids, -sign * sc))
            sc, ids = sc[order], ids[order]
            sc[n_valid:] = -sign * np.finfo(np.float32).max
            ids[n_valid:] = -1
            S[g, q], I[g, q] = sc, ids
    if nq > 3:
        I[:, 3, :] = -1                                                          # a query without any valid entry
        S[:, 3, :] = -sign * np.finfo(np.float32).max
    D, J = merge_topk(torch.from_numpy(S).cuda(), torch.from_numpy(I).cuda(), k_out, metric)
    D, J = D.cpu().numpy(), J.cpu().numpy()
    for q in range(nq):
        s, i = S[:, q].reshape(-1), I[:, q].reshape(-1)
        ok = i >= 0
        s, i = s[ok], i[ok]
        order = np.lexsort((i, -sign * s))[:k_out]
        m = len(order)
        np.testing.assert_array_equal(D[q, :m], s[order])
        if G * k <= 4096:
            np.testing.assert_array_equal(J[q, :m], i[order])       # the rank merge: ties in ascending id order, always
        else:
            # select + sort: WHICH of the entries tied with the k_out-th score are kept is not defined (the reference's argsort
            # and Faiss's heap do not define it either); every id must be a real entry with the reported score, none twice
            score_of = dict(zip(i.tolist(), s.tolist()))
            assert len(set(J[q, :m].tolist())) == m
            assert all(score_of[int(a)] == float(b) for a, b in zip(J[q, :m], D[q, :m]))
            inner = D[q, :m] != D[q, m - 1]
            np.testing.assert_array_equal(J[q, :m][inner], i[order][inner])
        assert (J[q, m:] == -1).all() and (D[q, m:] == -sign * np.finfo(np.float32).max).all()
