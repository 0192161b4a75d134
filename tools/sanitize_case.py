"""Small all-paths case for compute-sanitizer (run under ONE tool per gpurun call): both kernel variants, ragged
fragments, a relaunched plan, the verify pass and the no-allocation checksum, checked against each other on the GPU."""
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import numpy as np

import dissem_b200 as d

cuts = [0, 5, 262144 - 3, 262144 + 7, 600_001, 1_000_003]
size = cuts[-1]
for variant in (1, 0):
    with d.Engine([0, 0, 0, 0]) as eng:
        eng.set_option("variant", variant)
        for n in (0, 1, 2):
            eng.alloc(n, 11, size)
            eng.fill(n, 11, d.FILL_PATTERN, 7)
        eng.alloc(3, 11, size)
        eng.fill(3, 11, d.FILL_POISON)
        s = eng.compile([(i % 3, 3, 11, a, b - a, size) for i, (a, b) in enumerate(zip(cuts, cuts[1:]))])
        for _ in range(2):
            eng.wait(eng.launch(s))
        ref = eng.verify(0, 11)
        assert np.array_equal(eng.checksums(3, 11), ref) and np.array_equal(eng.verify(3, 11), ref)
        assert np.array_equal(eng.read(3, 11), eng.read(0, 11))
        sz, ptr = eng.info(3, 11)
        assert np.array_equal(d.checksum_noalloc(0, ptr, sz), ref)
        xs = d.plan_relay([d.Xfer(0, 1, 11, 0, 0, size, size)] , 3)
        eng.reset(1, 11)
        eng.wait(eng.submit(xs))
        assert np.array_equal(eng.checksums(1, 11), ref)
        eng.sched_free(s)
print("sanitize case ok")
