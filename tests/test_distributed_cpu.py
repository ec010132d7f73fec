"""CPU, world_size 2, gloo: the exchange logic of the row-sharded search (shard ownership, global ids,
one all-gather of per-shard lists, merge).  The per-shard search and the merge are injected oracle
functions here because the real ones are CUDA kernels; what is under test is the multi-process plumbing."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    import oracle
    from nexus_b200.sharded import all_gather_lists, shard_bounds
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(0)
        c = rng.standard_normal((1001, 32)).astype(np.float32)
        x = rng.standard_normal((13, 32)).astype(np.float32)
        k = 10
        lo, hi = shard_bounds(c.shape[0], world)[rank]
        s, i = oracle.blocked_ip_search(c[lo:hi], x, k)              # stands in for the local CUDA search
        i = np.where(i >= 0, i + lo, -1)                              # id offset = first global row of the shard
        Ds, Is = all_gather_lists(torch.from_numpy(s), torch.from_numpy(i), world)
        assert Ds.shape == (world, 13, k) and Is.dtype == torch.int64
        # merge (oracle stand-in for nxf_merge_topk): concatenate lists, re-select
        cat_s = Ds.permute(1, 0, 2).reshape(13, -1).numpy()
        cat_i = Is.permute(1, 0, 2).reshape(13, -1).numpy()
        order = np.lexsort((cat_i, -cat_s), axis=1)[:, :k]
        ms, mi = np.take_along_axis(cat_s, order, 1), np.take_along_axis(cat_i, order, 1)
        ref_s, ref_i = oracle.blocked_ip_search(c, x, k)
        rep = oracle.check_topk(ref_s, ref_i, ms, mi, c, x)
        assert rep.ok, rep.messages
        np.save(os.path.join(out_dir, f"ids_{rank}.npy"), mi)
    finally:
        dist.destroy_process_group()


def test_two_rank_shard_gather_merge(tmp_path):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    a, b = np.load(tmp_path / "ids_0.npy"), np.load(tmp_path / "ids_1.npy")
    np.testing.assert_array_equal(a, b)            # every rank ends with the same merged result


def _np_merge(Ds, Is, k, metric="ip"):
    """numpy stand-in for nxf_merge_topk: [G, Q, k] lists -> [Q, k], (score desc, id asc), -1 padding last."""
    import torch
    G, Q, kk = Ds.shape
    cat_s = Ds.permute(1, 0, 2).reshape(Q, -1).numpy()
    cat_i = Is.permute(1, 0, 2).reshape(Q, -1).numpy()
    order = np.lexsort((cat_i, -cat_s), axis=1)[:, :k]
    return torch.from_numpy(np.take_along_axis(cat_s, order, 1)), torch.from_numpy(np.take_along_axis(cat_i, order, 1))


def _worker_partitioned(rank, world, port, out_dir):
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch
    import torch.distributed as dist
    import oracle
    from nexus_b200.sharded import all_gather_lists, exchange_partitioned, shard_bounds
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        rng = np.random.default_rng(1)
        c = rng.standard_normal((1501, 32)).astype(np.float32)
        for nq in (13, 12, 1):                                        # ragged (last block padded), even, fewer queries than ranks
            x = rng.standard_normal((nq, 32)).astype(np.float32)
            k = 10
            lo, hi = shard_bounds(c.shape[0], world)[rank]
            s, i = oracle.blocked_ip_search(c[lo:hi], x, k)
            i = np.where(i >= 0, i + lo, -1)
            D, I = torch.from_numpy(s), torch.from_numpy(i)
            Dp, Ip = exchange_partitioned(D, I, world, k, "ip", None, _np_merge)
            Ds, Is = all_gather_lists(D, I, world)
            Dg, Ig = _np_merge(Ds, Is, k)
            assert Dp.shape == (nq, k) and Ip.dtype == torch.int64
            np.testing.assert_array_equal(Ip.numpy(), Ig.numpy())     # both exchanges give the same merged lists
            np.testing.assert_array_equal(Dp.numpy(), Dg.numpy())
            ref_s, ref_i = oracle.blocked_ip_search(c, x, k)
            rep = oracle.check_topk(ref_s, ref_i, Dp.numpy(), Ip.numpy(), c, x)
            assert rep.ok, rep.messages
            np.save(os.path.join(out_dir, f"pids_{nq}_{rank}.npy"), Ip.numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_query_partitioned_exchange_equals_all_gather(tmp_path, world):
    """all-to-all of per-shard lists by query block + merge + all-gather of the merged blocks == all-gather + merge."""
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker_partitioned, args=(world, port, str(tmp_path)), nprocs=world, join=True)
    for nq in (13, 12, 1):
        ref = np.load(tmp_path / f"pids_{nq}_0.npy")
        for r in range(1, world):
            np.testing.assert_array_equal(np.load(tmp_path / f"pids_{nq}_{r}.npy"), ref)
