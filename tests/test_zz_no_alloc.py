"""dissem_checksum_noalloc — the entry point for hosts that cannot allocate device memory (an inference
server whose pool already owns all HBM; a latency-critical path that cannot take cudaMalloc's locks).

It must give exactly what dissem_layer_verify gives (bit-exact: integer sums), on two calls in a row — the
second call finds the first call's leftovers in the shared built-in scratch and must not be affected — and it
must refuse layers beyond its stated limit.  Both entry points are also held against the CPU oracle's checksum of
the same seeded bytes, so the test does not rest on the library agreeing with itself.  This file sorts last on
purpose: a run that stops at the first failure has then already run every other test."""
import numpy as np
import pytest

import dissem_b200 as d
from oracle import dissem_oracle as o

pytestmark = pytest.mark.gpu


def test_noalloc_checksum_matches_verify_twice_in_a_row():
    with d.Engine([0, 0]) as eng:
        big = 9 * d.CHUNK_BYTES + 4321            # first call: ten chunks, the last one short
        small = 2 * d.CHUNK_BYTES + 16            # second call: fewer chunks, other contents
        eng.alloc(0, 1, big)
        eng.fill(0, 1, d.FILL_PATTERN, 0xABCDEF)
        eng.alloc(1, 2, small)
        eng.fill(1, 2, d.FILL_PATTERN, 42)
        eng.alloc(0, 3, big)
        eng.fill(0, 3, d.FILL_POISON)
        want = {k: eng.verify(n, k) for (n, k) in ((0, 1), (1, 2), (0, 3))}   # the existing entry point
        # ... which must itself agree with the CPU oracle on these bytes (pattern fill / poison are the oracle's too)
        assert np.array_equal(want[1], o.checksum(o.fill_bytes(o.FILL_PATTERN, 1, big, 0xABCDEF)))
        assert np.array_equal(want[2], o.checksum(o.fill_bytes(o.FILL_PATTERN, 2, small, 42)))
        assert np.array_equal(want[3], o.checksum(o.fill_bytes(o.FILL_POISON, 3, big, 0)))
        for (node, layer) in ((0, 1), (1, 2), (0, 3), (0, 1)):               # back to back, sizes and contents changing
            size, ptr = eng.info(node, layer)
            got = d.checksum_noalloc(0, ptr, size)
            assert got.shape == want[layer].shape
            assert np.array_equal(got, want[layer]), f"layer {layer}: no-allocation pass differs from dissem_layer_verify"
        # interleaving with the allocating entry point does not disturb either
        assert np.array_equal(eng.verify(1, 2), want[2])
        size, ptr = eng.info(1, 2)
        assert np.array_equal(d.checksum_noalloc(0, ptr, size), want[2])


def test_noalloc_on_a_second_gpu_uses_that_gpus_own_scratch():
    """Each GPU holds its own copy of the built-in scratch array and the library looks its address up per GPU: a
    buffer that lives on GPU 1 is checked with device=1, twice in a row, against the oracle and against GPU 0's
    result for the same bytes."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    size = 5 * d.CHUNK_BYTES + 777
    want = o.checksum(o.fill_bytes(o.FILL_PATTERN, 9, size, 0x5EED))
    with d.Engine([0, 1]) as eng:
        for node in (0, 1):
            eng.alloc(node, 9, size)
            eng.fill(node, 9, d.FILL_PATTERN, 0x5EED)
        assert np.array_equal(eng.verify(1, 9), want)
        for _ in range(2):
            for node in (1, 0, 1):
                sz, ptr = eng.info(node, 9)
                assert np.array_equal(d.checksum_noalloc(node, ptr, sz), want), f"device {node}"


def test_noalloc_rejects_sizes_beyond_its_scratch():
    too_big = (d.NOALLOC_MAX_CHUNKS + 1) * d.CHUNK_BYTES
    with pytest.raises(d.DissemError) as e:
        d.checksum_noalloc(0, 256, too_big)       # refused before the pointer is ever used
    assert e.value.code == -7 and "scratch holds" in str(e.value)
    with d.Engine([0]) as eng:
        eng.alloc(0, 0, 1000)
        size, ptr = eng.info(0, 0)
        with pytest.raises(d.DissemError) as e:
            d.lib().dissem_checksum_noalloc.restype  # noqa: B018  (binding exists)
            out = np.zeros((3, 2), dtype=np.uint64)
            rc = d.lib().dissem_checksum_noalloc(0, ptr, size, out.ctypes.data, 3, None)   # wrong chunk count
            if rc != 0:
                raise d.DissemError(rc, d.lib().dissem_last_error().decode())
        assert e.value.code == -1
        assert d.checksum_noalloc(0, ptr, 0).shape == (0, 2)                             # empty input: nothing to do
