"""Does one NVLink port carry more when its traffic is SPLIT between a pushed stream (stores issued by the source GPU) and
a pulled stream (loads issued by the destination GPU)?  A push tops out near 713 GB/s per direction, a pull near 788 GB/s
while the reverse direction is idle; the copy engine writes 777 GB/s.  ONE process over 2 GPUs, L layers of S bytes per
direction, device time of one launch (max over the two GPUs' streams, CUDA events), best of `reps`.

    python tools/p2p_mix_probe.py [--layers 8] [--size 1073741824] [--reps 3]
With --one CASE only that case runs once more after a warm-up (the ncu target: push | pull | mix | relay).
"""
import argparse
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d


def setup(eng, L, S, both):
    for l in range(L):
        eng.alloc(0, l, S)
        eng.fill(0, l, d.FILL_PATTERN, 0x5EED)
        eng.alloc(1, l, S)
    if both:
        for l in range(L, 2 * L):
            eng.alloc(1, l, S)
            eng.fill(1, l, d.FILL_PATTERN, 0x5EED)
            eng.alloc(0, l, S)


def case_xfers(name, L, S):
    """0->1 carries layers 0..L-1, 1->0 carries layers L..2L-1.  pull_frac of each direction's layers are pulled."""
    P = d.XFER_PULL
    fwd = lambda l, flag: d.Xfer(0, 1, l, flag, 0, S, S)
    rev = lambda l, flag: d.Xfer(1, 0, l, flag, 0, S, S)
    cases = {
        "0->1 push": ([fwd(l, 0) for l in range(L)], False),
        "0->1 pull": ([fwd(l, P) for l in range(L)], False),
        "0->1 half pushed, half pulled": ([fwd(l, P if l % 2 else 0) for l in range(L)], False),
        "0->1 one in four pulled": ([fwd(l, P if l % 4 == 0 else 0) for l in range(L)], False),
        "0->1 three in four pulled": ([fwd(l, 0 if l % 4 == 0 else P) for l in range(L)], False),
        "both directions, all pushed": ([fwd(l, 0) for l in range(L)] + [rev(L + l, 0) for l in range(L)], True),
        "both directions, all pulled": ([fwd(l, P) for l in range(L)] + [rev(L + l, P) for l in range(L)], True),
        "both directions, each half pushed half pulled": ([fwd(l, P if l % 2 else 0) for l in range(L)] +
                                                          [rev(L + l, P if l % 2 else 0) for l in range(L)], True),
        "both directions, GPU 0 executes everything (push out, pull in)": ([fwd(l, 0) for l in range(L)] +
                                                                           [rev(L + l, P) for l in range(L)], True),
    }
    return cases[name] if name else cases


def run_case(name, L, S, reps):
    xs, both = case_xfers(name, L, S)
    eng = d.Engine([0, 1])
    try:
        setup(eng, L, S, both)
        s = eng.compile(xs)
        eng.wait(eng.launch(s))
        best = min(eng.wait_detail(eng.launch(s))[0] for _ in range(reps))
        ok = bool((eng.checksums(1, 0) == eng.verify(0, 0)).all())
        per_dir = L * S
        print(json.dumps({"case": name, "layers_per_direction": L, "layer_bytes": S, "ms": best,
                          "GBps_per_direction": per_dir / best / 1e6, "verified_self_check": ok}), flush=True)
        eng.sched_free(s)
    finally:
        eng.close()


def run_one(which, L, S):
    """One launch of one kernel kind after a warm-up launch: the ncu target."""
    if which == "relay":
        # a forwarding round: GPU 1 forwards to GPU 0's second node what it received from GPU 0 (single process: rounds are
        # ordered by events; the kernel of interest is the one GPU 1 launches)
        eng = d.Engine([0, 1, 0])
        for l in range(L):
            eng.alloc(0, l, S)
            eng.fill(0, l, d.FILL_PATTERN, 0x5EED)
            eng.alloc(1, l, S)
            eng.alloc(2, l, S)
        base = [d.Xfer(0, n, l, 0, 0, S, S) for l in range(L) for n in (1, 2)]
        xs = d.plan_relay(base, 2)
    else:
        eng = d.Engine([0, 1])
        setup(eng, L, S, False)
        flag = {"push": lambda l: 0, "pull": lambda l: d.XFER_PULL, "mix": lambda l: d.XFER_PULL if l % 2 else 0}[which]
        xs = [d.Xfer(0, 1, l, flag(l), 0, S, S) for l in range(L)]
    s = eng.compile(xs)
    eng.wait(eng.launch(s))
    ms = eng.wait_detail(eng.launch(s))[0]
    print(json.dumps({"one": which, "ms": ms, "transfers": len(xs)}), flush=True)
    eng.sched_free(s)
    eng.close()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--size", type=int, default=1 << 30)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--one", default="")
    args = ap.parse_args()
    if torch.cuda.device_count() < 2:
        print(json.dumps({"error": "needs 2 GPUs"}))
        return
    if args.one:
        run_one(args.one, args.layers, args.size)
        return
    for name in case_xfers("", args.layers, args.size):
        run_case(name, args.layers, args.size, args.reps)


if __name__ == "__main__":
    main()
