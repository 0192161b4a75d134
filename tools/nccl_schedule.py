"""Internal comparison point only (north star: "An NCCL broadcast/send-recv version of the same schedule is kept only
as an internal comparison point"): the fan-out placement of tools/fanout_bench.py executed by NCCL under torchrun,
one rank per GPU —  (a) one grouped ncclSend/ncclRecv batch for the whole plan,  (b) one ncclBroadcast per layer on a
sub-communicator of {seed} U destinations.  Not part of the product path; nothing in libdissem.so calls NCCL.

    torchrun --nproc-per-node 8 tools/nccl_schedule.py [--layers 80] [--size 1711308800] [--one-seed]
"""
import argparse
import json
import os
import time

import torch
import torch.distributed as dist


def fanout_plan(n_nodes, n_layers, one_seed):
    plan = []
    for l in range(n_layers):
        seed = 0 if one_seed else (l + 4) % n_nodes
        dsts = [l % n_nodes, (l + 1) % n_nodes]
        if l % 3 >= 1:
            dsts.append((l + 2) % n_nodes)
        if l % 3 == 2:
            dsts.append((l + 3) % n_nodes)
        plan.append((l, seed, sorted(set(dsts) - {seed})))
    return plan


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=80)
    ap.add_argument("--size", type=int, default=1_711_308_800)
    ap.add_argument("--one-seed", action="store_true")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--bcast", action="store_true", help="also run the broadcast-per-layer rendition (did not finish on the 8-GPU box)")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    plan = fanout_plan(world, args.layers, args.one_seed)
    bufs = {}
    for l, seed, dsts in plan:
        if rank == seed:
            bufs[l] = torch.full((args.size,), (l * 7 + 1) % 251, dtype=torch.uint8, device="cuda")
        elif rank in dsts:
            bufs[l] = torch.zeros(args.size, dtype=torch.uint8, device="cuda")
    D = sum(len(dsts) for _l, _s, dsts in plan) * args.size
    tag = "one-seed" if args.one_seed else "balanced"

    def timed(fn):
        for _ in range(2):
            fn()
        torch.cuda.synchronize()
        dist.barrier()
        best = 1e9
        for _ in range(args.reps):
            torch.cuda.synchronize()
            dist.barrier()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            dt = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            best = min(best, dt.item())
        return best

    # (a) grouped send/recv of the whole plan
    def sendrecv():
        ops = []
        for l, seed, dsts in plan:
            if rank == seed:
                ops += [dist.P2POp(dist.isend, bufs[l], d) for d in dsts]
            elif rank in dsts:
                ops.append(dist.P2POp(dist.irecv, bufs[l], seed))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    t_sr = timed(sendrecv)
    ok = all(bool((bufs[l][:: max(1, args.size // 4096)] == (l * 7 + 1) % 251).all()) for l in bufs)
    if rank == 0:
        print(json.dumps({"case": f"{tag}/NCCL grouped send-recv", "n_gpus": world, "completion_ms_best": t_sr * 1e3,
                          "aggregate_GBps": D / t_sr / 1e9, "delivered_bytes": D, "verified": ok}), flush=True)

    if not args.bcast:
        dist.barrier()
        dist.destroy_process_group()
        return
    # (b) broadcast per layer on a sub-communicator
    groups = {}
    for l, seed, dsts in plan:
        key = tuple(sorted([seed] + dsts))
        if key not in groups:
            groups[key] = dist.new_group(list(key))      # collective over all ranks
    def bcast():
        works = []
        for l, seed, dsts in plan:
            key = tuple(sorted([seed] + dsts))
            if rank in key:
                works.append(dist.broadcast(bufs[l], src=seed, group=groups[key], async_op=True))
        for w in works:
            w.wait()

    t_bc = timed(bcast)
    if rank == 0:
        print(json.dumps({"case": f"{tag}/NCCL broadcast per layer", "n_gpus": world, "completion_ms_best": t_bc * 1e3,
                          "aggregate_GBps": D / t_bc / 1e9, "delivered_bytes": D, "sub_communicators": len(groups)}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
