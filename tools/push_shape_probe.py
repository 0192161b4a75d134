"""SM push over NVLink, ring-shape sweep (ONE process, 2 GPUs): bulk-store size (TMA ring stage bytes x stages) x CTAs per SM,
one direction and both directions at once, 8 x 256 MiB per direction; kernel time from CUDA events, best of 5.
Completes the shapes round 1 swept on the register path (vector width, threads, loads in flight).

    python tools/push_shape_probe.py
"""
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d

SIZE, NL = 256 << 20, 8
RINGS = {0: "16Kx6", 1: "32Kx3", 2: "8Kx12", 3: "16Kx4", 4: "8Kx6", 5: "4Kx12"}


def main():
    if torch.cuda.device_count() < 2:
        print(json.dumps({"error": "needs 2 GPUs"}))
        return
    eng = d.Engine([0, 1])
    for l in range(NL):
        eng.alloc(0, l, SIZE); eng.fill(0, l, d.FILL_PATTERN, 1); eng.alloc(1, l, SIZE)
        eng.alloc(1, 50 + l, SIZE); eng.fill(1, 50 + l, d.FILL_PATTERN, 2); eng.alloc(0, 50 + l, SIZE)
    uni = [(0, 1, l, 0, SIZE, SIZE) for l in range(NL)]
    bidi = uni + [(1, 0, 50 + l, 0, SIZE, SIZE) for l in range(NL)]
    eng.set_option("variant", 1)
    for name, xs in (("0->1", uni), ("0<->1", bidi)):
        for cfg, ring in RINGS.items():
            for ctas in (1, 2, 0):
                if ctas == 2 and cfg in (4, 5):
                    pass
                eng.set_option("tma_cfg", cfg)
                eng.set_option("ctas_per_sm", ctas)
                s = eng.compile(xs)
                for _ in range(2):
                    eng.wait(eng.launch(s))
                best = min(eng.wait_detail(eng.launch(s))[1] for _ in range(5))
                print(json.dumps({"case": name, "ring": ring, "ctas_per_sm": ctas or "occupancy", "kernel_ms": best,
                                  "GBps_per_direction": NL * SIZE / best / 1e6}), flush=True)
                eng.sched_free(s)
    ok = bool((eng.checksums(1, 0) == eng.verify(0, 0)).all())
    print(json.dumps({"self_check": ok}))
    eng.close()


if __name__ == "__main__":
    main()
