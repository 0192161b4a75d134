"""Copy-engine comparison point (not a product path): cudaMemcpyPeerAsync 0->1 alone, and 0->1 together with 1->0,
through torch.  Answers whether the copy engine keeps its one-directional rate when the reverse direction of the port
is busy, as SM loads do not (profiles/r01_summary.md).  One JSON line per case.   python tools/ce_bidir.py"""
import json
import time

import torch

N = 2 << 30


def run(pairs, reps=10):
    streams = [torch.cuda.Stream(device=src.device) for src, _ in pairs]
    def once():
        for (src, dst), st in zip(pairs, streams):
            with torch.cuda.stream(st):
                dst.copy_(src, non_blocking=True)
    for _ in range(3):
        once()
    for dev in (0, 1):
        torch.cuda.synchronize(dev)
    t0 = time.perf_counter()
    for _ in range(reps):
        once()
    for dev in (0, 1):
        torch.cuda.synchronize(dev)
    return (time.perf_counter() - t0) / reps


def main():
    if torch.cuda.device_count() < 2:
        print(json.dumps({"error": "needs 2 GPUs"}))
        return
    a0 = torch.empty(N, dtype=torch.uint8, device="cuda:0")
    b1 = torch.empty(N, dtype=torch.uint8, device="cuda:1")
    c1 = torch.empty(N, dtype=torch.uint8, device="cuda:1")
    d0 = torch.empty(N, dtype=torch.uint8, device="cuda:0")
    dt = run([(a0, b1)])
    print(json.dumps({"case": "copy engine 0->1 alone", "GBps_per_direction": N / dt / 1e9}), flush=True)
    dt = run([(a0, b1), (c1, d0)])
    print(json.dumps({"case": "copy engine 0->1 and 1->0 together", "GBps_per_direction": N / dt / 1e9}), flush=True)
    dt = run([(a0[: N // 2], b1[: N // 2]), (a0[N // 2:], b1[N // 2:])])
    print(json.dumps({"case": "copy engine 0->1, two streams", "GBps_per_direction": N / dt / 1e9}), flush=True)


if __name__ == "__main__":
    main()
