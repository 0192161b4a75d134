"""Disk tier (DiskLayer / `-s`, cmd/config.go:133-157, transport.go:351-367): a layer kept in a file is
streamed through pinned staging buffers into the destination's HBM with the checksum fused.  Bar: the destination holds
exactly the file's section, its fused table equals the CPU oracle's checksum of those bytes."""
import importlib
import json
import os
import subprocess

import numpy as np
import pytest

import dissem_b200 as d
from oracle import dissem_oracle as o
from roles_oracle import check_against_oracle, run_checked

pytestmark = pytest.mark.gpu
host = importlib.import_module("distributed-llm-dissemination_b200.host")
SEED = 0x5EED
STAGE = 16 << 20          # kStageBytes of the engine: pieces are cut at multiples of it in layer space


def _file_with(tmp_path, node, layer, data, lead=0):
    """A layer file as CreateDiskLayer names it; `lead` junk bytes first when the layer starts at a file offset."""
    path = o.disk_layer_path(str(tmp_path), node, layer)
    os.makedirs(os.path.dirname(path), exist_ok=True)
    with open(path, "wb") as f:
        f.write(b"\x77" * lead)
        f.write(data.tobytes())
    return path


@pytest.mark.parametrize("size", [0, 1, 15, 4097, d.CHUNK_BYTES + 3, STAGE - 1, STAGE, STAGE + 1, 2 * STAGE + 5 * d.CHUNK_BYTES + 7])
def test_whole_layer_from_file(tmp_path, size):
    want = o.fill_bytes(o.FILL_PATTERN, 9, size, SEED)
    path = _file_with(tmp_path, 0, 9, want)
    with d.Engine([0, 0]) as eng:
        eng.attach_file(0, 9, path, size)
        assert eng.file_size(0, 9) == size
        eng.alloc(1, 9, size)
        eng.fill(1, 9, d.FILL_POISON)
        for _ in range(2):                                   # a whole layer covers every chunk: relaunching is idempotent
            eng.wait(eng.send_file(0, 1, 9, 0, size, size))
            got = eng.read(1, 9)
            assert np.array_equal(got, o.read_disk_section(path, 0, size))
            assert np.array_equal(got, want)
            assert np.array_equal(eng.checksums(1, 9), o.checksum(want))
            assert np.array_equal(eng.verify(1, 9), o.checksum(want))
        eng.detach_file(0, 9)
        with pytest.raises(d.DissemError) as e:
            eng.file_size(0, 9)
        assert e.value.code == -4


def test_fragments_from_a_file_offset(tmp_path):
    """Mode-3 fragments [Offset, Offset+DataSize) (node.go:1604-1606) cut at odd bytes, from a layer that starts
    inside its file; one send per fragment, as Transport.Send issues them."""
    size = STAGE + 3 * d.CHUNK_BYTES + 1001
    want = o.fill_bytes(o.FILL_PATTERN, 4, size, SEED)
    lead = 4099
    path = _file_with(tmp_path, 2, 4, want, lead=lead)
    cuts = [0, 13, d.CHUNK_BYTES - 1, 5 * d.CHUNK_BYTES + 77, STAGE - 3, STAGE + 5, size]
    with d.Engine([0, 0, 0]) as eng:
        eng.attach_file(2, 4, path, size, file_off=lead)
        eng.alloc(0, 4, size)
        eng.fill(0, 4, d.FILL_POISON)
        eng.reset(0, 4)
        tickets = [eng.send_file(2, 0, 4, a, b - a, size) for a, b in zip(cuts[:-1], cuts[1:])]
        for t in tickets:
            eng.wait(t)
        assert np.array_equal(eng.read(0, 4), want)
        assert np.array_equal(eng.checksums(0, 4), o.checksum(want))
        assert eng.received(0, 4) == size
        # one fragment alone: only its bytes land, and its chunks hold the fragment's share of the sums
        eng.fill(0, 4, d.FILL_POISON)
        eng.reset(0, 4)
        a, b = cuts[2], cuts[3]
        eng.wait(eng.send_file(2, 0, 4, a, b - a, size))
        got = eng.read(0, 4)
        assert np.array_equal(got[a:b], o.read_disk_section(path, lead + a, b - a))
        assert (got[:a] == o.POISON_BYTE).all() and (got[b:] == o.POISON_BYTE).all()
        assert np.array_equal(eng.checksums(0, 4), o.checksum_fragment(want[a:b], a, size))


def test_node_loads_its_own_disk_copy_and_mixed_plan(tmp_path):
    """A node may be its own destination (disk -> its HBM); one submitted plan may mix disk and HBM sources."""
    size = 3 * d.CHUNK_BYTES + 321
    a = o.fill_bytes(o.FILL_PATTERN, 1, size, SEED)
    b = o.fill_bytes(o.FILL_PATTERN, 2, size, SEED)
    path = _file_with(tmp_path, 0, 1, a)
    with d.Engine([0, 0, 0]) as eng:
        eng.attach_file(0, 1, path, size)
        eng.alloc(0, 1, size)                       # node 0's own HBM copy of layer 1, still empty
        eng.fill(0, 1, d.FILL_POISON)
        eng.alloc(0, 2, size)
        eng.fill(0, 2, d.FILL_PATTERN, SEED)
        for n in (1, 2):
            eng.alloc(n, 1, size)
            eng.alloc(n, 2, size)
            eng.fill(n, 1, d.FILL_POISON)
            eng.fill(n, 2, d.FILL_POISON)
        plan = [d.Xfer(0, 0, 1, d.XFER_FILE, 0, size, size),        # load
                d.Xfer(0, 1, 1, d.XFER_FILE, 0, size, size),        # disk -> peer
                d.Xfer(0, 1, 2, 0, 0, size, size),                  # HBM -> peer
                d.Xfer(0, 2, 2, 0, 0, size, size)]
        eng.wait(eng.submit(plan))
        for node, layer, want in ((0, 1, a), (1, 1, a), (1, 2, b), (2, 2, b)):
            assert np.array_equal(eng.read(node, layer), want)
            assert np.array_equal(eng.checksums(node, layer), o.checksum(want))
        # now that node 0 holds layer 1 in HBM, the same transfer without the flag reads that copy
        eng.wait(eng.send(0, 2, 1, 0, size, size))
        assert np.array_equal(eng.read(2, 1), a)


def test_disk_tier_errors(tmp_path):
    size = 70_000
    path = _file_with(tmp_path, 0, 3, o.fill_bytes(o.FILL_PATTERN, 3, size, SEED))
    with d.Engine([0, 0]) as eng:
        with pytest.raises(d.DissemError) as e:
            eng.attach_file(0, 3, str(tmp_path / "missing.layer"), size)
        assert e.value.code == -11
        with pytest.raises(d.DissemError) as e:                 # the file is shorter than offset + size
            eng.attach_file(0, 3, path, size, file_off=1)
        assert e.value.code == -11 and "fewer" in str(e.value)
        with pytest.raises(d.DissemError) as e:                 # nothing attached yet
            eng.alloc(1, 3, size)
            eng.send_file(0, 1, 3, 0, size, size)
        assert e.value.code == -4
        eng.attach_file(0, 3, path, size)
        with pytest.raises(d.DissemError) as e:
            eng.attach_file(0, 3, path, size)
        assert e.value.code == -5
        with pytest.raises(d.DissemError) as e:                 # TotalSize disagrees with the file
            eng.send_file(0, 1, 3, 0, size - 1, size - 1)
        assert e.value.code == -7
        with pytest.raises(d.DissemError) as e:
            eng.send_file(0, 1, 3, 10, size, size)
        assert e.value.code == -7
        with pytest.raises(d.DissemError) as e:                 # streamed, never compiled
            eng.compile([d.Xfer(0, 1, 3, d.XFER_FILE, 0, size, size)])
        assert e.value.code == -8
        with pytest.raises(d.DissemError) as e:                 # rounds are for HBM-resident relays
            eng.submit([d.Xfer(0, 1, 3, d.XFER_FILE | (1 << 8), 0, size, size)])
        assert e.value.code == -8
        eng.wait(eng.send_file(0, 1, 3, 0, size, size))          # and the engine still works afterwards
        assert np.array_equal(eng.read(1, 3), o.read_disk_section(path, 0, size))


@pytest.mark.parametrize("mode,batched", [(0, False), (1, False), (1, True), (2, False), (3, False), (3, True)])
def test_distribution_from_disk_layers(tmp_path, mode, batched):
    """`-s`: every node's initial layers are files (CreateDiskLayer); the roles send them as DiskLayer sources and
    the receivers end with in-memory copies (assignmentSatisfied wants Location == InmemLayer, node.go:440)."""
    size = 300_001
    text = o.config_ring_fixture(4, size) if mode else o.config_identity_fixture(4, size)
    cfg = o.parse_config(text)
    # one file exists beforehand with other bytes: it must be used as it is (cmd/config.go:140 writes only if missing)
    leader = cfg.leader
    l0 = sorted(leader.layers)[0]
    odd = o.fill_bytes(o.FILL_PATTERN, 1000 + l0, size, SEED + 1)
    o.create_disk_layer(str(tmp_path), leader.id, l0, size, contents=odd)
    r = run_checked(text, mode, batched=batched, seed=SEED, time_scale=1_000_000, flags=3, timeout_ms=20000,
                    storage=str(tmp_path), overrides={(leader.id, l0): odd})
    assert r["oracle_checked"] > 0 and r["ready_equals_assignment"] and r["layers_checked"] > 0
    for n in cfg.nodes:
        for layer, (sz, _) in n.layers.items():
            p = o.disk_layer_path(str(tmp_path), n.id, layer)
            assert os.path.getsize(p) == sz
            got = o.read_disk_section(p, 0, sz)
            want = odd if (n.id, layer) == (leader.id, l0) else o.fill_bytes(o.FILL_PATTERN, layer, sz, SEED)
            assert np.array_equal(got, want)


def test_cli_storage_flag(tmp_path):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "distributed-llm-dissemination_b200", "dissem")
    if not os.path.exists(exe):
        pytest.skip("CLI not built")
    cfg = os.path.join(root, "tests", "data", "cli_config_4x8.json")
    for mode, extra in (("1", []), ("2", []), ("3", ["--pace"]), ("0", ["--batched"])):
        r = subprocess.run([exe, "-id", "0", "-f", cfg, "-m", mode, "-s", str(tmp_path), "-v", "--dump-tables"] + extra, capture_output=True, text=True, timeout=180)
        assert r.returncode == 0, (mode, r.stderr)
        assert f"storagePath: {tmp_path}" in r.stdout and "Time to deliver: " in r.stdout
        report = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        assert report["ready_equals_assignment"] and check_against_oracle(open(cfg).read(), report, seed=24301) > 0, mode
    conf = o.parse_config(open(cfg).read())
    for n in conf.nodes:
        for layer, (sz, _) in n.layers.items():
            assert os.path.getsize(o.disk_layer_path(str(tmp_path), n.id, layer)) == sz


@pytest.mark.parametrize("mode", [0, 3])
def test_paced_sends_honour_limit_rate(mode):
    """`pace`: a layer send limited to LimitRate B/s (Sources, cmd/config.go:111; mode 3: dataSize/minTime, node.go:1281)
    takes at least what writeWithLimit's token bucket allows (transport.go:407-424), and the bytes are still exact."""
    size, rate = 1 << 20, 20_000_000
    conf = json.loads(o.config_identity_fixture(4, size))
    for n in conf["Nodes"]:
        n["Sources"] = {k: rate for k in n["Sources"]}
    text = json.dumps(conf)
    floor_ms = o.limiter_write_times(size, rate)[-1][0] / 1e6            # 39.3 ms for one layer
    fast = run_checked(text, mode, seed=SEED, time_scale=1_000_000, flags=2, timeout_ms=20000)
    paced = run_checked(text, mode, seed=SEED, time_scale=1_000_000, flags=2, timeout_ms=20000, pace=1)
    assert fast["oracle_checked"] == paced["oracle_checked"] == 4 and paced["ready_equals_assignment"]
    if mode == 3:
        # the flow plan's own rate: size * ticks-per-second / t, never above the source's limit
        floor_ms = o.limiter_write_times(size, min(rate, size * 1_000_000 // paced["t"]))[-1][0] / 1e6 * 0.98
    assert paced["time_to_deliver_ms"] >= floor_ms
    assert paced["time_to_deliver_ms"] < 20 * floor_ms
    assert fast["time_to_deliver_ms"] < paced["time_to_deliver_ms"]
    assert paced["launches"] > fast["launches"]


def test_file_to_peer_gpu_and_disk_distribution_over_two_gpus(tmp_path):
    """The source node's GPU reads the staged pieces over PCIe and writes them into a peer GPU's HBM over NVLink."""
    if d.lib().dissem_device_count() < 2:
        pytest.skip("needs 2 GPUs")
    size = STAGE + 5 * d.CHUNK_BYTES + 11
    want = o.fill_bytes(o.FILL_PATTERN, 5, size, SEED)
    path = _file_with(tmp_path, 0, 5, want)
    with d.Engine([0, 1]) as eng:
        eng.attach_file(0, 5, path, size)
        eng.alloc(1, 5, size)
        eng.fill(1, 5, d.FILL_POISON)
        eng.wait(eng.submit([d.Xfer(0, 1, 5, d.XFER_FILE, 0, 1000, size), d.Xfer(0, 1, 5, d.XFER_FILE, 1000, size - 1000, size)]))
        assert np.array_equal(eng.read(1, 5), want)
        assert np.array_equal(eng.checksums(1, 5), o.checksum(want))
        assert np.array_equal(eng.verify(1, 5), o.checksum(want))
    # files of nodes on different GPUs in one plan: streamed concurrently, one thread per GPU
    other = o.fill_bytes(o.FILL_PATTERN, 6, size, SEED)
    path6 = _file_with(tmp_path, 1, 6, other)
    with d.Engine([0, 1]) as eng:
        eng.attach_file(0, 5, path, size)
        eng.attach_file(1, 6, path6, size)
        for node in (0, 1):
            eng.alloc(node, 5, size)
            eng.alloc(node, 6, size)
        plan = [d.Xfer(0, 1, 5, d.XFER_FILE, 0, size, size), d.Xfer(1, 0, 6, d.XFER_FILE, 0, size, size),
                d.Xfer(0, 0, 5, d.XFER_FILE, 0, size, size), d.Xfer(1, 1, 6, d.XFER_FILE, 0, size, size)]
        for _ in range(2):
            eng.wait(eng.submit(plan))
            for node in (0, 1):
                assert np.array_equal(eng.read(node, 5), want) and np.array_equal(eng.read(node, 6), other)
                assert np.array_equal(eng.checksums(node, 5), o.checksum(want))
                assert np.array_equal(eng.checksums(node, 6), o.checksum(other))
        with pytest.raises(d.DissemError) as e:                  # an error inside a streaming thread reaches the caller
            os.truncate(path6, size // 2)
            eng.wait(eng.submit(plan))
        assert e.value.code == -11 and "ends at" in str(e.value)
    text = o.config_ring_fixture(4, 300_001)
    for mode, batched in ((1, False), (3, True)):
        store = tmp_path / f"m{mode}"
        r = run_checked(text, mode, node_device=[i % 2 for i in range(5)], batched=batched, seed=SEED,
                        time_scale=1_000_000, flags=3, timeout_ms=20000, storage=str(store))
        assert r["oracle_checked"] > 0 and r["ready_equals_assignment"]


def test_disk_layer_read_once_then_forwarded(tmp_path):
    """dissem_plan_optimize on a disk layer with three destinations: one file read, the rest forwarded from HBM; the
    engine starts the forwards only after the disk transfer has landed (also when they run on another GPU)."""
    n_gpu = max(1, min(d.lib().dissem_device_count(), 2))
    size = STAGE + 9 * d.CHUNK_BYTES + 5
    want = o.fill_bytes(o.FILL_PATTERN, 7, size, SEED)
    path = _file_with(tmp_path, 0, 7, want)
    gpus = [0, 1 % n_gpu, 0, 1 % n_gpu]
    with d.Engine(gpus) as eng:
        eng.attach_file(0, 7, path, size)
        for n in (1, 2, 3):
            eng.alloc(n, 7, size)
        plan, _ = d.plan_optimize([d.Xfer(0, n, 7, d.XFER_FILE, 0, size, size) for n in (1, 2, 3)], gpus)
        assert sum(1 for x in plan if x.flags & d.XFER_FILE) == 1
        for rep in range(3):
            for n in (1, 2, 3):
                eng.fill(n, 7, d.FILL_POISON)
            eng.wait(eng.submit(plan))
            for n in (1, 2, 3):
                assert np.array_equal(eng.read(n, 7), want), (rep, n)
                assert np.array_equal(eng.checksums(n, 7), o.checksum(want)), (rep, n)


def test_batched_distribution_from_disk_with_fanout(tmp_path):
    """The batched leader hands a plan with disk sources to dissem_plan_optimize: layers assigned to several nodes."""
    size = 300_001
    conf = json.loads(o.config_identity_fixture(4, size))
    conf["Assignment"] = {"1": {"1": {}, "2": {}}, "2": {"1": {}, "2": {}}, "3": {"1": {}, "3": {}}, "4": {"4": {}}}
    text = json.dumps(conf)
    for batched in (True, False):
        r = run_checked(text, 0, batched=batched, seed=SEED, timeout_ms=20000, storage=str(tmp_path / str(batched)))
        assert r["oracle_checked"] == 7 and r["ready_equals_assignment"] and r["layers_checked"] == 7


def test_diskspeed_tool(tmp_path):
    """diskspeed/main.go: `-path file` prints the file size, the time to load it and MiB/s; here the load ends in HBM."""
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = os.path.join(root, "distributed-llm-dissemination_b200", "diskspeed")
    if not os.path.exists(exe):
        pytest.skip("diskspeed not built")
    assert subprocess.run([exe], capture_output=True, text=True, timeout=60).stdout == "usage: -path [file path]\n"
    size = STAGE + 12345
    path = _file_with(tmp_path, 0, 0, o.fill_bytes(o.FILL_PATTERN, 0, size, SEED))
    r = subprocess.run([exe, "-path", path], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    lines = r.stdout.strip().splitlines()
    assert lines[0] == f"File size: {size}" and lines[1].startswith("Time to load: ") and lines[2].startswith("Throughput: ")
    assert float(lines[2].split()[1]) > 0 and lines[2].endswith("MiB/s")
    bad = subprocess.run([exe, "-path", str(tmp_path / "nope")], capture_output=True, text=True, timeout=60)
    assert bad.returncode != 0 and "failed to get fileInfo" in bad.stderr
