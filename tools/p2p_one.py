"""One push stream GPU0 -> GPU1 (4 x 256 MiB per launch, 4 launches) in one process: the smallest program that shows the
NVLink traffic of the copy kernel; meant to be run plain and then under ncu (single process, not multi-rank)."""
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import dissem_b200 as d

SIZE, NL = 256 << 20, 4
pull = len(sys.argv) > 1 and sys.argv[1] == "pull"
with d.Engine([0, 1]) as eng:
    eng.set_option("pull", 1 if pull else 0)
    if len(sys.argv) > 2:
        eng.set_option("variant", 0)
        eng.set_option("reg_cfg", int(sys.argv[2]))
    for l in range(NL):
        eng.alloc(0, l, SIZE)
        eng.fill(0, l, d.FILL_PATTERN, 1)
        eng.alloc(1, l, SIZE)
    s = eng.compile([(0, 1, l, 0, SIZE, SIZE) for l in range(NL)])
    ts = [eng.wait_detail(eng.launch(s))[1] for _ in range(4)]
    ok = bool((eng.checksums(1, 0) == eng.verify(0, 0)).all())
    print(json.dumps({"mode": "pull" if pull else "push", "reg_cfg": (int(sys.argv[2]) if len(sys.argv) > 2 else "tma"), "bytes_per_launch": NL * SIZE, "kernel_ms": ts,
                      "GBps_best": NL * SIZE / min(ts) / 1e6, "verified": ok}))
