"""Internal comparison point only (north star: "An NCCL broadcast/send-recv version of the same schedule is kept only
as an internal comparison point"): a fan-out placement executed by NCCL under torchrun, one rank per GPU —
  (a) one grouped ncclSend/ncclRecv batch for the whole plan,
  (b) one ncclBroadcast per layer on a sub-communicator of {seed} U destinations (communicators split off the world
      communicator once, outside the clock).
Not part of the product path; nothing in libdissem.so calls NCCL.  bench.py imports `compare` for its north_star block.

    torchrun --nproc-per-node 8 tools/nccl_schedule.py [--layers 80] [--size 1711308800] [--one-seed] [--bcast]
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


def _timed(fn, reps):
    """Best wall time of `reps` runs, max over ranks; one untimed run first."""
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        fn()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], device="cuda", dtype=torch.float64)
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        best = min(best, dt.item())
    return best


def compare(plan, size, rank, reps=2, do_bcast=True, groups_cache=None):
    """plan: [(layer, seed rank, [destination ranks])].  Allocates its own torch buffers (call it when the engine's
    layers have been freed), runs both renditions, frees the buffers.  Returns {"nccl_sendrecv_ms", "nccl_broadcast_ms",
    "sub_communicators", "nccl_verified"} on every rank."""
    bufs = {}
    for l, seed, dsts in plan:
        if rank == seed:
            bufs[l] = torch.full((size,), (l * 7 + 1) % 251, dtype=torch.uint8, device="cuda")
        elif rank in dsts:
            bufs[l] = torch.zeros(size, dtype=torch.uint8, device="cuda")
    out = {}

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

    out["nccl_sendrecv_ms"] = _timed(sendrecv, reps) * 1e3
    stride = max(1, size // 4096)
    ok = all(bool((bufs[l][::stride] == (l * 7 + 1) % 251).all()) for l in bufs)
    if do_bcast:
        for l, seed, dsts in plan:              # receivers start from zeros again
            if rank in dsts:
                bufs[l].zero_()
        # one communicator per distinct rank set, made once (ncclCommSplit when the world group was created with a
        # device id).  new_group is collective over ALL ranks and must be called in the same order everywhere.
        groups = groups_cache if groups_cache is not None else {}
        for l, seed, dsts in plan:
            key = tuple(sorted([seed] + dsts))
            if key not in groups:
                groups[key] = dist.new_group(list(key))

        def bcast():
            # Every rank walks the plan in the same order and chains its broadcasts (wait() orders the next call's
            # stream behind this one): collectives on different communicators that share GPUs must be issued in one
            # global order, or the cooperative kernels can wait for each other in a cycle.
            for l, seed, dsts in plan:
                key = tuple(sorted([seed] + dsts))
                if rank in key:
                    dist.broadcast(bufs[l], src=seed, group=groups[key], async_op=True).wait()

        out["nccl_broadcast_ms"] = _timed(bcast, reps) * 1e3
        out["sub_communicators"] = len(groups)
        ok = ok and all(bool((bufs[l][::stride] == (l * 7 + 1) % 251).all()) for l in bufs)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    out["nccl_verified"] = bool(flag.item())
    bufs.clear()
    torch.cuda.empty_cache()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=80)
    ap.add_argument("--size", type=int, default=1_711_308_800)
    ap.add_argument("--one-seed", action="store_true")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--bcast", action="store_true", help="also run the broadcast-per-layer rendition")
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    plan = fanout_plan(world, args.layers, args.one_seed)
    res = compare(plan, args.size, rank, reps=args.reps, do_bcast=args.bcast)
    if rank == 0:
        D = sum(len(dsts) for _l, _s, dsts in plan) * args.size
        res.update({"case": ("one-seed" if args.one_seed else "balanced") + "/NCCL", "n_gpus": world, "delivered_bytes": D})
        print(json.dumps(res), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
