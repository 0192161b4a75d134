"""The one-seed / balanced Llama-3-70B fan-out placement under torchrun, ONE PROCESS PER GPU (rank r owns node r):
unicast with direction choice versus relay chains whose rounds are separated by a barrier between the ranks, versus the
same chains with every round launched at once and the hops ordered on the device (DISSEM_XFER_WAIT_SRC).

    torchrun --nproc-per-node 8 tools/fanout_bench_mp.py [--layers 80] [--size 1711308800] [--one-seed] [--pieces 16]
"""
import argparse
import importlib
import json
import os
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch
import torch.distributed as dist

import dissem_b200 as d

multi = importlib.import_module("distributed-llm-dissemination_b200.multi")


def fanout_plan(n_nodes, n_layers, size, one_seed):
    holders, xfers = {}, []
    for l in range(n_layers):
        seed = 0 if one_seed else (l + 4) % n_nodes
        holders[l] = seed
        dsts = [l % n_nodes, (l + 1) % n_nodes]
        if l % 3 >= 1:
            dsts.append((l + 2) % n_nodes)
        if l % 3 == 2:
            dsts.append((l + 3) % n_nodes)
        for dst in sorted(set(dsts)):
            if dst != seed:
                xfers.append(d.Xfer(seed, dst, l, 0, 0, size, size))
    return holders, xfers


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=80)
    ap.add_argument("--size", type=int, default=1_711_308_800)
    ap.add_argument("--one-seed", action="store_true")
    ap.add_argument("--pieces", type=int, default=16)
    ap.add_argument("--reps", type=int, default=3)
    args = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        torch.cuda.synchronize()
        dist.barrier()

    owner = lambda node: node
    holders, base = fanout_plan(world, args.layers, args.size, args.one_seed)
    eng = d.Engine(multi.device_of_node(world, owner, rank, local))
    for l, seed in holders.items():
        if seed == rank:
            eng.alloc(rank, l, args.size)
            eng.fill(rank, l, d.FILL_PATTERN, 0x5EED)
    for x in base:
        if x.dst_node == rank and not eng.has(rank, x.layer):
            eng.alloc(rank, x.layer, args.size)
            eng.fill(rank, x.layer, d.FILL_POISON)
    tag = "one-seed" if args.one_seed else "balanced"
    D = sum(x.size for x in base)
    store = multi.SharedStore(eng, dist, world, rank, owner)
    relay = d.plan_directions(d.plan_relay(base, args.pieces), list(range(world)))
    cases = (("unicast + direction choice", d.plan_directions(base, list(range(world))), "rounds"),
             (f"relay chains, {args.pieces} pieces + direction choice, barrier between rounds", relay, "rounds"),
             (f"relay chains, {args.pieces} pieces + direction choice, hops ordered on the device (DISSEM_XFER_WAIT_SRC)", relay, "flags"))
    for name, xs, how in cases:
        store.share(xs)
        if how == "flags":
            rs = multi.FlagRelaySchedule(eng, xs, owner, rank)
            run = lambda: rs.run(barrier)[0]
        else:
            rs = multi.RoundedSchedule(eng, xs)
            run = lambda: rs.run(barrier)
        run()
        barrier()
        best, best_dev = 1e9, 1e9
        for _ in range(args.reps):
            barrier()
            t0 = time.perf_counter()
            dev_ms = run()
            barrier()
            dt = torch.tensor([time.perf_counter() - t0, dev_ms / 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
            best = min(best, dt[0].item())
            best_dev = min(best_dev, dt[1].item())
        ok = eng.get_option("wait_timeouts") == 0
        for x in base:
            if x.dst_node == rank:
                ok = ok and bool((eng.checksums(rank, x.layer) == eng.verify(rank, x.layer)).all()) and eng.checksums(rank, x.layer).any()
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if rank == 0:
            print(json.dumps({"case": f"{tag}/{name}", "processes": world, "rounds": 1 + max(d.xfer_round(x) for x in xs), "layers": args.layers,
                              "layer_bytes": args.size, "delivered_bytes": D, "completion_ms_best": best * 1e3,
                              "device_ms_best_max_over_ranks": best_dev * 1e3, "aggregate_GBps": D / best / 1e9,
                              "self_check": bool(flag.item())}), flush=True)
        rs.free()
        for x in base:                       # fresh tables and poison for the next case
            if x.dst_node == rank:
                eng.fill(rank, x.layer, d.FILL_POISON)
                eng.reset(rank, x.layer)
        barrier()
    eng.close()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
