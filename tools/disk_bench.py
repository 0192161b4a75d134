"""Disk tier throughput: one layer file -> dissem_send_file -> HBM, wall clock around send + wait,
for several reader-thread counts, beside (a) the file read alone into pinned memory and (b) the same bytes already
in pinned host memory (dissem_layer_attach_host), which is the PCIe bound.  The file was just written, so it is read
from the page cache: this measures the staging pipeline, not a storage device.  One JSON line per case.

    python tools/disk_bench.py [--size 2147483648] [--dir /tmp] [--threads 1,2,4,8,16] [--reps 3]
"""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])

import dissem_b200 as d


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=2 << 30)
    ap.add_argument("--dir", default="/tmp")
    ap.add_argument("--threads", default="1,2,4,8,16")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--gpus", type=int, default=8, help="also load one file per GPU concurrently on up to this many GPUs")
    args = ap.parse_args()
    size = args.size
    path = os.path.join(args.dir, f"dissem_disk_bench_{os.getpid()}.layer")
    with d.Engine([0, 0]) as eng:
        # contents from the fill kernel, written out through pinned memory
        eng.alloc(0, 0, size)
        eng.fill(0, 0, d.FILL_PATTERN, 0x5EED)
        ref = eng.verify(0, 0)
        hb = d.HostBuffer(size)
        try:
            d.lib().dissem_layer_read(eng._ctx, 0, 0, hb.ptr, 0, size)
            with open(path, "wb") as f:
                f.write(memoryview(hb.array))
            eng.free(0, 0)
            eng.alloc(1, 0, size)
            # (b) PCIe bound: the bytes are already in pinned host memory
            eng.attach_host(0, 0, hb.ptr, size)
            best = 1e9
            for _ in range(args.reps):
                t0 = time.perf_counter()
                eng.wait(eng.send(0, 1, 0, 0, size, size))
                best = min(best, time.perf_counter() - t0)
            ok = bool((eng.checksums(1, 0) == ref).all())
            print(json.dumps({"case": "pinned host source (attach_host)", "bytes": size, "best_s": best, "GBps": size / best / 1e9, "verified": ok}), flush=True)
            eng.free(0, 0)
            # (a) the read alone
            fd = os.open(path, os.O_RDONLY)
            t0 = time.perf_counter()
            pos = 0
            mv = memoryview(hb.array)
            while pos < size:
                pos += os.preadv(fd, [mv[pos:pos + (16 << 20)]], pos)
            dt = time.perf_counter() - t0
            os.close(fd)
            print(json.dumps({"case": "pread alone, 1 thread, 16 MiB calls into pinned memory", "bytes": size, "best_s": dt, "GBps": size / dt / 1e9}), flush=True)
        finally:
            hb.free()
        eng.attach_file(0, 0, path, size)
        for th in [int(t) for t in args.threads.split(",")]:
            eng.set_option("file_threads", th)
            best = 1e9
            for _ in range(args.reps):
                eng.fill(1, 0, d.FILL_POISON)
                eng.read(1, 0, 0, 16)                 # drain the fill
                t0 = time.perf_counter()
                eng.wait(eng.send_file(0, 1, 0, 0, size, size))
                best = min(best, time.perf_counter() - t0)
            ok = bool((eng.checksums(1, 0) == ref).all() and (eng.verify(1, 0) == ref).all())
            print(json.dumps({"case": "file -> staging ring -> HBM", "file_threads": th, "bytes": size, "best_s": best,
                              "GBps": size / best / 1e9, "verified": ok, "host_cores": os.cpu_count()}), flush=True)
    os.unlink(path)
    # several GPUs: one file per GPU, each loaded into a node of its own GPU, all in one submitted plan
    n_gpu = d.lib().dissem_device_count()
    if args.gpus > 1 and n_gpu >= 2:
        n = min(args.gpus, n_gpu)
        per = size // 2
        paths = []
        with d.Engine(list(range(n))) as eng:
            hb = d.HostBuffer(per)
            try:
                eng.alloc(0, 100, per)
                eng.fill(0, 100, d.FILL_PATTERN, 0x5EED)
                ref = eng.verify(0, 100)
                d.lib().dissem_layer_read(eng._ctx, 0, 100, hb.ptr, 0, per)
                for g in range(n):
                    p = os.path.join(args.dir, f"dissem_disk_bench_{os.getpid()}_{g}.layer")
                    with open(p, "wb") as f:
                        f.write(memoryview(hb.array))
                    paths.append(p)
                eng.free(0, 100)
            finally:
                hb.free()
            for g in range(n):
                eng.attach_file(g, 100, paths[g], per)
                eng.alloc(g, 100, per)
            for k in sorted({1, 2, n}):
                plan = [d.Xfer(g, g, 100, d.XFER_FILE, 0, per, per) for g in range(k)]
                best = 1e9
                for _ in range(args.reps):
                    t0 = time.perf_counter()
                    eng.wait(eng.submit(plan))
                    best = min(best, time.perf_counter() - t0)
                ok = all(bool((eng.checksums(g, 100) == ref).all()) for g in range(k))
                print(json.dumps({"case": "one file per GPU, loaded concurrently", "gpus_loading": k, "bytes_each": per,
                                  "best_s": best, "aggregate_GBps": k * per / best / 1e9, "verified": ok,
                                  "host_cores": os.cpu_count()}), flush=True)
        for p in paths:
            os.unlink(p)


if __name__ == "__main__":
    main()
