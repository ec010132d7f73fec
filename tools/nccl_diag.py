"""Transport diagnosis for the exchange step (run under torchrun): what NCCL uses between the ranks of this box, how long one
all_gather of the C2 exchange size takes, and what a plain peer copy achieves.  Prints one JSON line on rank 0."""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    import torch.distributed as dist
    world, rank, local = int(os.environ["WORLD_SIZE"]), int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    out = {"world": world}
    for nbytes in (8, 5_584_000, 80_000_000):
        mine = torch.zeros(nbytes // 8, dtype=torch.int64, device=dev)
        allb = torch.zeros(world * (nbytes // 8), dtype=torch.int64, device=dev)
        for _ in range(5):
            dist.all_gather_into_tensor(allb, mine)
        torch.cuda.synchronize()
        dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            dist.all_gather_into_tensor(allb, mine)
        e1.record()
        torch.cuda.synchronize()
        out[f"all_gather_{nbytes}B_us"] = e0.elapsed_time(e1) / 20 * 1e3
        t0 = time.perf_counter()
        for _ in range(20):
            dist.all_gather_into_tensor(allb, mine)
        out[f"all_gather_{nbytes}B_host_issue_us"] = (time.perf_counter() - t0) / 20 * 1e6
        torch.cuda.synchronize()
    if rank == 0 and torch.cuda.device_count() > 1:
        out["can_access_peer_0_1"] = bool(torch.cuda.can_device_access_peer(0, 1))
        a = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda:0")
        b = torch.zeros(64 << 20, dtype=torch.uint8, device="cuda:1")
        for _ in range(3):
            b.copy_(a)
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            b.copy_(a, non_blocking=True)
        e1.record()
        torch.cuda.synchronize(0); torch.cuda.synchronize(1)
        out["peer_copy_64MB_GBps"] = (64 << 20) * 10 / (e0.elapsed_time(e1) * 1e-3) / 1e9
    dist.barrier()
    if rank == 0:
        print(json.dumps(out))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
