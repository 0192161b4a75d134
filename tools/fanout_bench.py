"""BASELINE config 4 in ONE process over all visible GPUs: Llama-3-70B-shaped layers, each assigned to 2-4 GPUs
(fan-out cycling 2/3/4), plus the one-seed variant.  Compares, on the same placement and with device-resident
layers:  unicast push | unicast with per-transfer direction choice | NVLS multicast (default path).
Prints one JSON line per case.  No oracle, no CPU path; correctness is checked on the GPU afterwards
(fused table at every destination == independent re-read == re-read of the source).

    python tools/fanout_bench.py [--layers 80] [--size 1711308800] [--one-seed] [--reps 5]
"""
import argparse
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d

LINK_GBPS = 770.0       # measured peer copy per direction (B200_PROFILING.md); nominal 900


def fanout_plan(n_nodes, n_layers, size, one_seed):
    """layer l -> {l, l+1, (l+2 if l%3>=1), (l+3 if l%3==2)} mod N, seeded on GPU (l+4) mod N."""
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


def t_star(xfers, n, multicast):
    """max over GPUs of max(egress, ingress) / link; a multicast layer counts once on its source's egress."""
    eg, ing = [0] * n, [0] * n
    seen = set()
    for x in xfers:
        ing[x.dst_node] += x.size
        if multicast:
            if (x.src_node, x.layer) in seen:
                continue
            seen.add((x.src_node, x.layer))
        eg[x.src_node] += x.size
    return max(max(e, i) for e, i in zip(eg, ing)) / (LINK_GBPS * 1e9), eg, ing


def run_case(name, n, holders, xfers, size, vmm, mcast, directions, reps, variant=1, relay=0, optimize=False, flags=False):
    eng = d.Engine(list(range(n)))
    try:
        eng.set_option("variant", variant)
        eng.set_option("alloc_vmm", 1 if vmm else 0)
        eng.set_option("mcast", mcast)
        if flags:
            eng.set_option("wait_src", 1)        # rounds ordered by device-side waits alone: counters must start at zero
        for l, seed in holders.items():
            eng.alloc(seed, l, size)
            eng.fill(seed, l, d.FILL_PATTERN, 0x5EED)
        for x in xfers:
            if not eng.has(x.dst_node, x.layer):
                eng.alloc(x.dst_node, x.layer, size)
                eng.fill(x.dst_node, x.layer, d.FILL_POISON)
        xs = d.plan_relay(xfers, relay) if relay else xfers
        xs = d.plan_directions(xs, list(range(n))) if directions else xs
        if optimize:
            xs, relay = d.plan_optimize(xfers, list(range(n)))
            name += f" -> {'relay, %d pieces' % relay if relay else 'unicast'}"
        s = eng.compile(xs)
        groups = eng.get_option("mcast_groups")
        dsts = sorted({(x.dst_node, x.layer) for x in xfers})

        def once():
            if flags:
                for (node, layer) in dsts:
                    eng.reset(node, layer)
            return eng.wait_detail(eng.launch(s))[0]

        for _ in range(2):
            once()
        times = [once() for _ in range(reps)]
        if flags:
            assert eng.get_option("wait_timeouts") == 0
        ok = True
        for x in xfers[:: max(1, len(xfers) // 24)]:       # sample of destinations
            ref = eng.verify(x.src_node, x.layer)
            ok = ok and bool((eng.checksums(x.dst_node, x.layer) == ref).all() and (eng.verify(x.dst_node, x.layer) == ref).all())
        best, avg = min(times), sum(times) / len(times)
        D = sum(x.size for x in xfers)
        ts, eg, ing = t_star(xs if relay else xfers, n, multicast=(groups > 0))
        print(json.dumps({"case": name, "n_gpus": n, "layers": len(holders), "layer_bytes": size, "deliveries": len(xfers),
                          "delivered_bytes": D, "mcast_groups": groups, "transfers_pulled": sum(1 for x in xs if x.flags & 1), "rounds": 1 + max(d.xfer_round(x) for x in xs),
                          "completion_ms_best": best, "completion_ms_avg": avg, "aggregate_GBps": D / best / 1e6,
                          "t_star_ms": ts * 1e3, "frac_of_link_roofline": ts * 1e3 / best, "busiest_egress": max(eg),
                          "busiest_ingress": max(ing), "verified": ok}), flush=True)
        eng.sched_free(s)
    finally:
        eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=80)
    ap.add_argument("--size", type=int, default=1_711_308_800)
    ap.add_argument("--one-seed", action="store_true")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--cases", default="unicast,auto,mcast,default")
    args = ap.parse_args()
    n = min(torch.cuda.device_count(), 8)
    if n < 3:
        print(json.dumps({"error": "needs 3+ GPUs"}))
        return
    holders, xfers = fanout_plan(n, args.layers, args.size, args.one_seed)
    tag = "one-seed" if args.one_seed else "balanced"
    cases = args.cases.split(",")
    if "unicast" in cases:
        run_case(f"{tag}/unicast push", n, holders, xfers, args.size, vmm=False, mcast=0, directions=False, reps=args.reps)
    if "auto" in cases:
        run_case(f"{tag}/unicast auto direction", n, holders, xfers, args.size, vmm=False, mcast=0, directions=True, reps=args.reps)
    if "mcast" in cases:
        run_case(f"{tag}/NVLS multicast (always)", n, holders, xfers, args.size, vmm=True, mcast=3, directions=False, reps=args.reps)
    if "default" in cases:
        run_case(f"{tag}/default: cost-model multicast + auto direction", n, holders, xfers, args.size, vmm=True, mcast=1,
                 directions=True, reps=args.reps)
    if "mcast_reg" in cases:
        run_case(f"{tag}/NVLS multicast (register kernel)", n, holders, xfers, args.size, vmm=True, mcast=3, directions=False,
                 reps=args.reps, variant=0)
    for c in cases:
        if c.startswith("relay"):            # relay16, relay32p (p = with direction choice)
            pieces = int("".join(ch for ch in c if ch.isdigit()) or 16)
            fl = c.endswith("pf")                 # relay8pf: direction choice + hops ordered by device-side waits (no round events)
            run_case(f"{tag}/relay chains, {pieces} pieces" + (" + direction choice" if c.endswith("p") or fl else "") +
                     (", hops ordered on the device" if fl else ""), n, holders,
                     xfers, args.size, vmm=False, mcast=0, directions=c.endswith("p") or fl, reps=args.reps, relay=pieces, flags=fl)
    if "optimized" in cases:
        run_case(f"{tag}/dissem_plan_optimize", n, holders, xfers, args.size, vmm=False, mcast=0, directions=False,
                 reps=args.reps, optimize=True)
    if "mcast_auto" in cases:
        run_case(f"{tag}/NVLS multicast + auto direction for the rest", n, holders, xfers, args.size, vmm=True, mcast=1,
                 directions=True, reps=args.reps)


if __name__ == "__main__":
    main()
