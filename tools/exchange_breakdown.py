"""Where a step of the row-sharded search goes outside the main kernel (run under torchrun, one rank per GPU):

    torchrun --nproc-per-node 2 tools/exchange_breakdown.py [--config c2] [--iters 20]

CUDA events around the three phases of DistributedFlatIndex.search -- the shard's own search (packed output), the one NCCL
all-gather, the merge kernel -- plus the library's own event timing of the main kernel inside the first phase."""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c2")
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    import torch
    import torch.distributed as dist
    from nexus_b200 import DistributedFlatIndex, merge_packed, shard_bounds
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    Q, N, D, k, store, norm, desc = bench.CONFIGS[a.config]
    lo, hi = shard_bounds(N, world)[rank]
    didx = DistributedFlatIndex(D, "ip", store, device=local)
    didx.local.reserve(hi - lo)
    didx.local.set_id_offset(lo)
    bench.build_shard(torch, didx.local, lo, hi, D, store, norm, dev, seed=bench.CORPUS_SEED[a.config])
    didx._ntotal_global, didx._offsets = N, [b[0] for b in shard_bounds(N, world)]
    q = bench.make_queries(torch, Q, D, norm, dev, seed=bench.QUERY_SEED[a.config])
    mine, lists = didx._buffers(Q, k, dev)
    for _ in range(3):
        didx.search(q, k)
    torch.cuda.synchronize()
    ev = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(a.iters)]
    didx.local.set_profiling(True)
    dist.barrier()
    torch.cuda.synchronize()
    for it in range(a.iters):
        ev[it][0].record()
        didx.local.search_packed(q, k, out=mine)
        ev[it][1].record()
        dist.all_gather_into_tensor(lists, mine)
        ev[it][2].record()
        merge_packed(lists, didx._offsets, k, "ip")
        ev[it][3].record()
    torch.cuda.synchronize()
    st = didx.local.stats()
    ph = [sum(ev[it][j].elapsed_time(ev[it][j + 1]) for it in range(a.iters)) / a.iters for j in range(3)]
    tot = sum(ev[it][0].elapsed_time(ev[it][3]) for it in range(a.iters)) / a.iters
    kern = st["main_kernel_ms"] / a.iters
    out = torch.tensor([ph[0], ph[1], ph[2], tot, kern], device=dev, dtype=torch.float64)
    dist.all_reduce(out, op=dist.ReduceOp.MAX)
    if rank == 0:
        o = out.tolist()
        print(json.dumps({"config": a.config, "n_gpus": world, "ms_search_packed": o[0], "ms_main_kernel_inside": o[4], "ms_search_outside_kernel": o[0] - o[4],
                          "ms_all_gather": o[1], "ms_merge_packed": o[2], "ms_step": o[3], "note": "max over ranks of per-rank means, CUDA events on the current stream"}))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
