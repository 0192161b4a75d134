"""Host-to-device copy rate per GPU, one rank at a time and all ranks at once (torchrun, one rank per GPU): is the end-to-end
leg of bench.py at 8 GPUs (24 GB/s per GPU) bound by the box (host memory / PCIe root) or by the engine's upload path?
Plain torch copies from pinned memory, no engine involved.

    torchrun --nproc-per-node 8 tools/h2d_probe.py [--size 268435456] [--reps 10]
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=256 << 20)
    ap.add_argument("--reps", type=int, default=10)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    host = torch.empty(args.size, dtype=torch.uint8).pin_memory()
    host.fill_(rank + 1)
    dev = torch.empty(args.size, dtype=torch.uint8, device="cuda")

    def copies():
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(args.reps):
            dev.copy_(host, non_blocking=True)
        torch.cuda.synchronize()
        return args.size * args.reps / (time.perf_counter() - t0) / 1e9

    copies()
    alone = torch.zeros(world, dtype=torch.float64, device="cuda")
    for r in range(world):                       # one rank at a time
        dist.barrier()
        if r == rank:
            alone[r] = copies()
        dist.barrier()
    dist.all_reduce(alone)
    together = torch.zeros(world, dtype=torch.float64, device="cuda")
    dist.barrier()
    together[rank] = copies()
    dist.barrier()
    dist.all_reduce(together)
    if rank == 0:
        print(json.dumps({"ranks": world, "bytes_per_copy": args.size, "GBps_alone_per_rank": [round(v, 1) for v in alone.tolist()],
                          "GBps_all_at_once_per_rank": [round(v, 1) for v in together.tolist()],
                          "GBps_all_at_once_total": round(float(together.sum()), 1)}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
