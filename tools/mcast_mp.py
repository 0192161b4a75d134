"""NVLS multicast with ONE PROCESS PER GPU (torchrun): rank 0 seeds every layer, all other ranks receive it.  Layers are
shareable VMM allocations; buffer and multicast handles travel between the processes as file descriptors.
Compares the multicast stream with unicast push on the same buffers.

    torchrun --nproc-per-node 4 tools/mcast_mp.py [--layers 8] [--size 268435456]
"""
import argparse
import importlib
import json
import os
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch
import torch.distributed as dist

import dissem_b200 as d

multi = importlib.import_module("distributed-llm-dissemination_b200.multi")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--size", type=int, default=256 << 20)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()

    fx = multi.FdExchange(rank, os.environ.get("MASTER_PORT", "0"))
    barrier()
    eng = d.Engine([local if n == rank else -1 for n in range(world)])
    if not eng.get_option("mcast_supported") and world > 0:
        pass                                        # single local device: the query needs 2; the create call will tell
    eng.set_option("alloc_vmm", 2)
    dsts = list(range(1, world))
    for l in range(args.layers):
        eng.alloc(rank, l, args.size)
        eng.fill(rank, l, d.FILL_PATTERN if rank == 0 else d.FILL_POISON, 0x5EED)
    barrier()
    # destination buffers -> rank 0
    if rank != 0:
        for l in range(args.layers):
            fx.send(0, f"buf:{rank}:{l}", eng.export_fd(rank, l))
    else:
        for r in dsts:
            for l in range(args.layers):
                eng.import_fd(r, l, args.size, fx.recv(f"buf:{r}:{l}"))
    barrier()
    # multicast objects: create on rank 0, join elsewhere, then everybody binds
    ok_mc, why = True, ""
    try:
        if rank == 0:
            for l in range(args.layers):
                fd = eng.mcast_create(l, 0, dsts)
                for r in dsts:
                    fx.send(r, f"mc:{l}", fd, keep=True)
                os.close(fd)
        else:
            for l in range(args.layers):
                eng.mcast_join(l, 0, dsts, rank, fx.recv(f"mc:{l}"))
    except Exception as e:                          # noqa: BLE001
        ok_mc, why = False, repr(e)
    flag = torch.tensor([1 if ok_mc else 0], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if not flag.item():
        if rank == 0 or why:
            print(json.dumps({"rank": rank, "error": "multicast set-up failed", "why": why}), flush=True)
        eng.close(); fx.close(); dist.destroy_process_group()
        return
    barrier()
    for l in range(args.layers):
        eng.mcast_bind(l, 0, dsts)
    barrier()
    xfers = [d.Xfer(0, r, l, 0, 0, args.size, args.size) for l in range(args.layers) for r in dsts]
    D = len(xfers) * args.size
    results = {}
    for name, mc in (("unicast push", 0), ("NVLS multicast", 2 if world == 2 else 3)):
        best, groups = None, 0
        if rank == 0:
            eng.set_option("mcast", mc)
            s = eng.compile(xfers)
            groups = eng.get_option("mcast_groups")
            for _ in range(2):
                eng.wait(eng.launch(s))
        barrier()
        if rank == 0:
            best = min(eng.wait_detail(eng.launch(s))[0] for _ in range(args.reps))
            eng.sched_free(s)
        barrier()
        ok = True
        if rank != 0:
            for l in range(args.layers):
                ok = ok and bool((eng.checksums(rank, l) == eng.verify(rank, l)).all()) and eng.checksums(rank, l).any()
        f2 = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(f2, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(json.dumps({"case": f"one process per GPU / {name}", "processes": world, "layers": args.layers, "layer_bytes": args.size,
                              "mcast_groups_total": groups, "completion_ms_best": best, "aggregate_GBps": D / best / 1e6,
                              "source_egress_GBps": (args.layers * args.size if mc else D) / best / 1e6, "verified": bool(f2.item())}), flush=True)
        for l in range(args.layers):                # fresh tables / poison for the next case
            if rank != 0:
                eng.fill(rank, l, d.FILL_POISON)
                eng.reset(rank, l)
        barrier()
    eng.close()
    fx.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
