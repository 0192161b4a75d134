"""NVLS multicast store stream, kernel shape sweep (ONE process over all visible GPUs): GPU 0 seeds L layers for k other GPUs.
For every shape — register path (threads x loads in flight), TMA path (consumer warps that issue multimem.st), CTAs per SM —
prints the rate at which the stream leaves the seed and lands at each destination port, next to unicast push / pull of the
same bytes to ONE destination.  The numbers behind the schedule compiler's cost model (dissem_engine.cu) come from here.

    python tools/mcast_probe.py [--layers 4] [--size 1073741824] [--dsts 1,3] [--reps 3] [--quick]
"""
import argparse
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d


def run(n, layers, size, dsts, mcast, variant, reg_cfg, tma_cfg, ctas, reps, pull=False, tag=""):
    eng = d.Engine(list(range(n)))
    try:
        eng.set_option("alloc_vmm", 1)
        eng.set_option("mcast", mcast)
        eng.set_option("variant", variant)
        eng.set_option("mcast_variant", 1)          # the sweep chooses the kernel itself
        eng.set_option("reg_cfg", reg_cfg)
        eng.set_option("tma_cfg", tma_cfg)
        eng.set_option("ctas_per_sm", ctas)
        for l in range(layers):
            eng.alloc(0, l, size)
            eng.fill(0, l, d.FILL_PATTERN, 0x5EED)
            for dn in dsts:
                eng.alloc(dn, l, size)
                eng.fill(dn, l, d.FILL_POISON)
        xs = [d.Xfer(0, dn, l, d.XFER_PULL if pull else 0, 0, size, size) for l in range(layers) for dn in dsts]
        s = eng.compile(xs)
        groups = eng.get_option("mcast_groups")
        eng.wait(eng.launch(s))
        best = min(eng.wait_detail(eng.launch(s))[0] for _ in range(reps))
        ref = eng.verify(0, 0)
        ok = all(bool((eng.checksums(dn, 0) == ref).all()) and bool((eng.verify(dn, layers - 1) == eng.verify(0, layers - 1)).all()) for dn in dsts)
        sent = layers * size * (1 if groups else len(dsts))
        print(json.dumps({"case": tag, "n_gpus": n, "dsts": len(dsts), "layers": layers, "layer_bytes": size, "mcast_groups": groups,
                          "variant": "tma" if variant else "register", "reg_cfg": reg_cfg, "tma_cfg": tma_cfg, "ctas_per_sm": ctas,
                          "ms": best, "seed_egress_GBps": sent / best / 1e6, "landing_GBps_per_port": layers * size / best / 1e6,
                          "verified": ok}), flush=True)
        eng.sched_free(s)
    finally:
        eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=4)
    ap.add_argument("--size", type=int, default=1 << 30)
    ap.add_argument("--dsts", default="")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--quick", action="store_true")
    ap.add_argument("--one", action="store_true", help="one multicast case only (register 512x4, all other GPUs as destinations): the ncu target")
    args = ap.parse_args()
    n = min(torch.cuda.device_count(), 8)
    if n < 2:
        print(json.dumps({"error": "needs 2+ GPUs"}))
        return
    ks = [int(v) for v in args.dsts.split(",") if v] or sorted({1, n - 1})
    L, S, R = args.layers, args.size, args.reps
    if args.one:
        k = n - 1
        run(n, L, S, list(range(1, k + 1)), 2 if k == 1 else 3, 0, 1, 0, 0, 1, tag=f"multicast x{k}, register cfg 1")
        return
    # unicast references: one destination
    run(n, L, S, [1], 0, 1, 1, 0, 0, R, tag="unicast push, TMA")
    run(n, L, S, [1], 0, 0, 1, 0, 0, R, tag="unicast push, register 512x4")
    run(n, L, S, [1], 0, 1, 1, 0, 0, R, pull=True, tag="unicast pull, TMA")
    for k in ks:
        dsts = list(range(1, k + 1))
        mc = 2 if k == 1 else 3
        reg = [(1, 0), (3, 0), (0, 0), (4, 0), (6, 0)] if not args.quick else [(1, 0), (3, 0)]
        for (cfg, ctas) in reg:
            run(n, L, S, dsts, mc, 0, cfg, 0, ctas, R, tag=f"multicast x{k}, register cfg {cfg}")
        for ctas in ([1, 3] if not args.quick else []):
            run(n, L, S, dsts, mc, 0, 1, 0, ctas, R, tag=f"multicast x{k}, register cfg 1, {ctas} CTAs/SM")
        for cfg in ([0, 6, 7, 8] if not args.quick else [0, 7]):
            run(n, L, S, dsts, mc, 1, 1, cfg, 0, R, tag=f"multicast x{k}, TMA cfg {cfg}")
        if k > 1:
            run(n, L, S, dsts, 0, 1, 1, 0, 0, R, tag=f"unicast push x{k} (no multicast)")


if __name__ == "__main__":
    main()
