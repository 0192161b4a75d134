"""NVLink probe (2+ GPUs, one process): push-kernel bandwidth GPU0 -> GPU1, both directions at once, fan-out,
against the copy engine (cudaMemcpyPeer via torch) as an internal comparison.  Prints one JSON line per case."""
import json
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d

SIZE = 256 << 20
NL = 8


def timed(eng, sched, reps=10):
    for _ in range(3):
        eng.wait(eng.launch(sched))
    ts = [eng.wait_detail(eng.launch(sched)) for _ in range(reps)]
    return min(t[1] for t in ts), sum(t[1] for t in ts) / len(ts)


def main():
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        print(json.dumps({"error": "needs 2 GPUs"}))
        return
    nodes = list(range(min(ngpu, 8)))
    eng2 = d.Engine(nodes)
    for l in range(NL):
        eng2.alloc(0, l, SIZE); eng2.fill(0, l, d.FILL_PATTERN, 1)
        eng2.alloc(1, l, SIZE)
        eng2.alloc(1, 50 + l, SIZE); eng2.fill(1, 50 + l, d.FILL_PATTERN, 2)
        eng2.alloc(0, 50 + l, SIZE)
    uni = [(0, 1, l, 0, SIZE, SIZE) for l in range(NL)]
    bidi = uni + [(1, 0, 50 + l, 0, SIZE, SIZE) for l in range(NL)]
    for name, xs in (("0->1", uni), ("0<->1", bidi)):
        for variant, pull, ctas, cfg in ((0, 0, 0, 1), (0, 0, 1, 1), (0, 1, 0, 1), (0, 1, 0, 3), (0, 1, 1, 1),
                                         (1, 0, 0, 1), (1, 0, 1, 1), (1, 1, 0, 1)):
            eng2.set_option("variant", variant)
            eng2.set_option("pull", pull)
            eng2.set_option("ctas_per_sm", ctas)
            eng2.set_option("reg_cfg", cfg)
            s = eng2.compile(xs)
            best, avg = timed(eng2, s)
            per_dir = NL * SIZE
            print(json.dumps({"case": name, "variant": "tma" if variant else "reg", "mode": "pull" if pull else "push",
                              "ctas_per_sm": ctas, "reg_cfg": cfg, "kernel_ms_best": best, "kernel_ms_avg": avg,
                              "GBps_per_direction_best": per_dir / best / 1e6, "GBps_per_direction_avg": per_dir / avg / 1e6}), flush=True)
            eng2.sched_free(s)
    eng2.set_option("variant", 0); eng2.set_option("pull", 0); eng2.set_option("ctas_per_sm", 0); eng2.set_option("reg_cfg", 1)
    # verify one landing buffer
    ok = bool((eng2.verify(1, 0) == eng2.verify(0, 0)).all() and (eng2.checksums(1, 0) == eng2.verify(0, 0)).all())
    print(json.dumps({"verified": ok, "received": eng2.received(1, 0)}))
    # copy engine comparison
    a = torch.empty(NL * SIZE, dtype=torch.uint8, device="cuda:0")
    b = torch.empty(NL * SIZE, dtype=torch.uint8, device="cuda:1")
    for _ in range(3):
        b.copy_(a)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    t0 = time.perf_counter()
    for _ in range(10):
        b.copy_(a)
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    dt = (time.perf_counter() - t0) / 10
    print(json.dumps({"case": "cudaMemcpyPeer 0->1 (copy engine, torch)", "GBps": NL * SIZE / dt / 1e9}))
    if ngpu >= 4:
        # fan-in: 1,2,3 -> 0 simultaneously (ingress-bound at node 0)
        eng3 = d.Engine(nodes)
        xs = []
        for src in (1, 2, 3):
            for l in range(4):
                lid = src * 10 + l
                eng3.alloc(src, lid, SIZE); eng3.fill(src, lid, d.FILL_PATTERN, 3)
                eng3.alloc(0, lid, SIZE)
                xs.append((src, 0, lid, 0, SIZE, SIZE))
        s = eng3.compile(xs)
        best, avg = timed(eng3, s)
        print(json.dumps({"case": "fan-in 1,2,3->0", "kernel_ms_best": best, "ingress_GBps_best": 12 * SIZE / best / 1e6}))
        eng3.close()
    eng2.close()


if __name__ == "__main__":
    main()
