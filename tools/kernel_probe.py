"""Single-GPU probe of the secondary kernels: read-only verify pass (K5) and pattern fill (K6) on one
Llama-3-70B-sized layer, wall-clock around the C-ABI call (includes the table's D2H for verify)."""
import json
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d

SIZE = 1_711_308_800
with d.Engine([0]) as eng:
    eng.alloc(0, 0, SIZE)
    for variant in (1, 0):
        eng.set_option("variant", variant)
        eng.fill(0, 0, d.FILL_PATTERN, 1)
        eng.verify(0, 0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(10):
            eng.verify(0, 0)
        dt = (time.perf_counter() - t0) / 10
        print(json.dumps({"kernel": "verify (K5)", "variant": "tma" if variant else "register", "bytes": SIZE,
                          "ms_wall": dt * 1e3, "GBps_read": SIZE / dt / 1e9}))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(10):
        eng.fill(0, 0, d.FILL_PATTERN, 2)
    eng.read(0, 0, 0, 16)
    dt = (time.perf_counter() - t0) / 10
    print(json.dumps({"kernel": "fill_pattern (K6)", "bytes": SIZE, "ms_wall": dt * 1e3, "GBps_written": SIZE / dt / 1e9}))
    size, ptr = eng.info(0, 0)
    d.checksum_noalloc(0, ptr, size)
    t0 = time.perf_counter()
    for _ in range(10):
        d.checksum_noalloc(0, ptr, size)
    dt = (time.perf_counter() - t0) / 10
    print(json.dumps({"kernel": "dissem_checksum_noalloc", "bytes": SIZE, "ms_wall": dt * 1e3, "GBps_read": SIZE / dt / 1e9}))
