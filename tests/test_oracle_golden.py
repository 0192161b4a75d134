"""CPU tests that pin the oracle (oracle/) against everything the reference offers for this path:
its own test fixtures (distributor/node_test.go), its README and conf/config.json inputs, and the
hand-derived flow-plan KAT.  Also: numpy oracle == C oracle, checksum properties."""
import json
import os

import numpy as np
import pytest

from oracle import c_oracle as co
from oracle import dissem_oracle as o

REF = "/root/reference"
GOLD = os.path.join(os.path.dirname(__file__), "golden")


# ---- fixtures of the reference's own tests -------------------------------------------------
def test_identity_fixture_mode0():
    """node_test.go:174-218 (TestSimpleDistribution/inmem, /tcp): 4 transfers 0->k of layer k,
    and the ready assignment equals the input assignment (node_test.go:133-134)."""
    cfg = o.parse_config(o.config_identity_fixture())
    xs = o.plan_mode0(cfg)
    assert [(x.src, x.dst, x.layer, x.off, x.size, x.total) for x in xs] == \
        [(0, k, k, 0, 1, 1) for k in range(1, 5)]
    store = o.apply_transfers(o.initial_store(cfg), xs)
    assert o.assignment_satisfied(cfg, store)
    assert cfg.assignment == {k: [k] for k in range(1, 5)}


def test_ring_fixture_mode1():
    """node_test.go:219-245 (tcp_retransmission) with the ring placement of :45-72: node 1 holds 4,
    node 2 holds 1, node 3 holds 2, node 4 holds 3; every needed layer has owners {leader, one peer}."""
    cfg = o.parse_config(o.config_ring_fixture())
    assert {n.id: sorted(n.layers) for n in cfg.nodes} == {0: [1, 2, 3, 4], 1: [4], 2: [1], 3: [2], 4: [3]}
    cand = o.mode1_owner_candidates(cfg)
    assert cand == {(1, 1): [0, 2], (2, 2): [0, 3], (3, 3): [0, 4], (4, 4): [0, 1]}
    for seed in range(8):
        for policy in ("random", "balanced"):
            xs = o.plan_mode1(cfg, seed=seed, policy=policy)
            assert len(xs) == 4
            for x in xs:
                assert x.src in cand[(x.dst, x.layer)] and (x.off, x.size, x.total) == (0, 1, 1)
            assert o.assignment_satisfied(cfg, o.apply_transfers(o.initial_store(cfg), xs))
    # the seeded pick really exercises both owners
    srcs = {x.src for s in range(16) for x in o.plan_mode1(cfg, seed=s)}
    assert 0 in srcs and len(srcs) > 1


def test_readme_config_both_modes():
    """readme.md:17-64 (flat schema; fed to HEAD it would yield zero layers)."""
    cfg = o.parse_config(o.config_readme())
    assert cfg.schema == "flat" and cfg.layer_size == 1048576
    assert {n.id: sorted(n.layers) for n in cfg.nodes} == {0: [1, 3], 1: [1], 2: [], 3: [3]}
    m0 = o.plan_mode0(cfg)
    assert [(x.src, x.dst, x.layer, x.size) for x in m0] == [(0, 2, 1, 1048576), (0, 2, 3, 1048576)]
    assert o.mode1_owner_candidates(cfg) == {(2, 1): [0, 1], (2, 3): [0, 3]}
    assert o.delivered_bytes(m0) == 2_097_152


@pytest.mark.skipif(not os.path.exists(f"{REF}/conf/config.json"), reason="reference tree not present")
def test_conf_json_rebuild_matches_reference_file():
    real = o.parse_config(open(f"{REF}/conf/config.json").read())
    assert real == o.parse_config(o.config_conf_json())
    assert real.schema == "nested"


@pytest.mark.skipif(not os.path.exists(f"{REF}/readme.md"), reason="reference tree not present")
def test_readme_json_verbatim_from_reference():
    text = open(f"{REF}/readme.md").read()
    block = text.split("```json")[1].split("```")[0]
    real = o.parse_config(block)
    mine = o.parse_config(o.config_readme())
    assert real.assignment == mine.assignment and real.layer_size == mine.layer_size
    assert [(n.id, n.is_leader, n.layers) for n in real.nodes] == [(n.id, n.is_leader, n.layers) for n in mine.nodes]


# ---- flow planner KAT ----------------------------------------------------------------------
def test_flow_kat_conf_json():
    """Hand-derived from flow.go's capacity model: t = 60 s; required flow
    87 445 534 144; every sender ships <= 209 715 200 * 60; fragments tile each layer."""
    cfg = o.parse_config(o.config_conf_json())
    t, xs, self_jobs = o.plan_mode3(cfg)
    assert t == 60 and self_jobs == []
    size = 10_930_691_768
    assert sum(x.size for x in xs) == 8 * size == 87_445_534_144
    per_sender = {}
    for x in xs:
        assert x.dst == 7 and x.total == size and x.size > 0
        per_sender[x.src] = per_sender.get(x.src, 0) + x.size
    assert set(per_sender) <= set(range(7)) and max(per_sender.values()) <= 209_715_200 * 60
    for layer in range(8):
        frs = sorted((x.off, x.size) for x in xs if x.layer == layer)
        pos = 0
        for off, sz in frs:
            assert off == pos
            pos += sz
        assert pos == size
    # t = 59 is infeasible in the same model (7 * 209 715 200 * 59 < required)
    assert 7 * 209_715_200 * 59 < 87_445_534_144 <= 7 * 209_715_200 * 60
    gold = json.load(open(os.path.join(GOLD, "flow_conf_json.json")))
    assert gold["t"] == t
    assert gold["xfers"] == [[x.src, x.dst, x.layer, x.off, x.size, x.total] for x in xs]


def test_flow_time_scale_and_single_dest():
    cfg = o.parse_config(o.config_stage_ring(4, 8, 1 << 20, ring=False))
    # drop the leader's own stage from the assignment? no: leader already holds it -> self job
    t, xs, self_jobs = o.plan_mode3(cfg, time_scale=1_000_000)
    assert len(self_jobs) == 2 and t >= 1
    store = o.apply_transfers(o.initial_store(cfg), xs)
    for n, ls in cfg.assignment.items():
        for l in ls:
            assert np.array_equal(store[(n, l)], o.fill_bytes(o.FILL_PATTERN, l, 1 << 20))


# ---- content + checksum ---------------------------------------------------------------------
@pytest.mark.parametrize("size", [0, 1, 3, 4, 5, 8, 15, 16, 17, 4095, 262143, 262144, 262145, 786432 + 13])
def test_numpy_oracle_equals_c_oracle(size):
    for kind in (o.FILL_ZERO, o.FILL_PATTERN, o.FILL_POISON):
        a = o.fill_bytes(kind, 5, size, 0x5EED)
        b = co.fill_bytes(kind, 5, size, 0x5EED)
        assert np.array_equal(a, b)
        assert np.array_equal(o.checksum(a), co.checksum(a))
    assert o.checksum(o.fill_bytes(o.FILL_ZERO, 0, size)).sum() == 0      # reference layers: A = B = 0


def test_checksum_golden_vectors():
    gold = json.load(open(os.path.join(GOLD, "checksum_kat.json")))
    for case in gold["cases"]:
        data = o.fill_bytes(case["kind"], case["layer"], case["size"], case["seed"])
        cs = o.checksum(data)
        assert [[int(a), int(b)] for a, b in cs.tolist()] == case["csum"]
        assert o.digest(cs) == case["digest"]
        assert int(data[:16].tobytes().hex() == case["head_hex"]) == 1


def test_checksum_additive_over_fragments_and_position_sensitive():
    size = 3 * o.CHUNK_BYTES + 1234
    data = o.fill_bytes(o.FILL_PATTERN, 9, size, 1)
    cuts = [0, 5, 262144 - 3, 262144 + 7, 600001, size]
    tot = np.zeros_like(o.checksum(data))
    with np.errstate(over="ignore"):
        for a, b in zip(cuts, cuts[1:]):
            tot += o.checksum_fragment(data[a:b], a, size)
    assert np.array_equal(tot, o.checksum(data))
    moved = data.copy()
    moved[4096:8192], moved[8192:12288] = data[8192:12288].copy(), data[4096:8192].copy()
    cm, cd = o.checksum(moved), o.checksum(data)
    assert cm[0, 0] == cd[0, 0] and cm[0, 1] != cd[0, 1]     # A blind to the swap, B sees it


def test_transfer_semantics_poison_and_offsets():
    cfg = o.parse_config(o.config_readme())
    store = o.initial_store(cfg, o.FILL_PATTERN, 3)
    half = 1048576 // 2
    xs = [o.Xfer(0, 2, 1, 0, half, 1048576)]
    o.apply_transfers(store, xs)
    got = store[(2, 1)]
    assert np.array_equal(got[:half], store[(0, 1)][:half]) and (got[half:] == o.POISON_BYTE).all()
    o.apply_transfers(store, [o.Xfer(1, 2, 1, half, half, 1048576)])
    assert np.array_equal(store[(2, 1)], store[(0, 1)])
    with pytest.raises(ValueError):
        o.apply_transfers(store, [o.Xfer(0, 2, 1, half, 1048576, 1048576)])


def test_baseline_configs_delivered_bytes():
    """BASELINE.md §3 table."""
    c2 = o.parse_config(o.config_stage_ring(8, 32, o.S2))
    assert o.delivered_bytes(o.plan_mode1(c2, seed=0)) == 28 * o.S2 == 1_879_048_192
    bal = o.plan_mode1(c2, policy="balanced")
    assert o.delivered_bytes(bal) == 28 * o.S2
    c3 = o.parse_config(o.config_stage_ring(8, 32, o.S8B, ring=False))
    assert o.delivered_bytes(o.plan_mode0(c3)) == 12_214_272_000
    c4 = o.parse_config(o.config_fanout())
    x4 = o.plan_mode1(c4)
    assert len(x4) == 27 * 2 + 27 * 3 + 26 * 4 == 239
    with pytest.raises(ValueError):
        o.plan_mode0(c4)


def test_tcp_baseline_delivers_bytes():
    cfg = o.parse_config(o.config_stage_ring(4, 8, 300_000))
    xs = o.plan_mode1(cfg, seed=2)
    layers = {l: co.fill_bytes(1, l, 300_000, 7) for l in range(8)}
    sec, moved = co.tcp_run(4, 0, xs, [layers[x.layer] for x in xs], base_port=19080, verify=True)
    assert moved == o.delivered_bytes(xs) and sec > 0
    # fragments (mode-3 style) are placed at their offsets
    frs = [o.Xfer(0, 1, 2, 0, 100_001, 300_000), o.Xfer(0, 1, 2, 100_001, 199_999, 300_000)]
    sec, moved = co.tcp_run(4, 0, frs, [layers[2], layers[2]], base_port=19180, verify=True)
    assert moved == 300_000


def test_disk_layer_restatement(tmp_path):
    """CreateDiskLayer (cmd/config.go:133-157): <storage>/layers/<node>/<layer>.layer, zeros, written only when missing;
    a DiskLayer goes on the wire as the file section [Offset, Offset+DataSize) (transport.go:356-364)."""
    p = o.create_disk_layer(str(tmp_path), 3, 11, 5000)
    assert p == str(tmp_path / "layers" / "3" / "11.layer")
    assert not o.read_disk_section(p, 0, 5000).any()
    marked = o.fill_bytes(o.FILL_PATTERN, 11, 5000, 7)
    with open(p, "r+b") as f:
        f.write(marked.tobytes())
    assert o.create_disk_layer(str(tmp_path), 3, 11, 5000) == p            # exists: left as it is
    assert np.array_equal(o.read_disk_section(p, 0, 5000), marked)
    assert np.array_equal(o.read_disk_section(p, 123, 1000), marked[123:1123])
    assert o.create_disk_layer(str(tmp_path), 3, 12, -4) and o.read_disk_section(o.disk_layer_path(str(tmp_path), 3, 12), 0, 10).size == 0


def test_tcp_restatement_disk_source_and_limit_rate(tmp_path):
    """The loopback-TCP restatement with (a) a DiskLayer source (sendLayerMsg copies the file section, transport.go:351-367)
    and (b) writeWithLimit's token bucket (transport.go:407-424): the paced send takes what limiter_write_times says."""
    size = 3 * 262144 + 999
    a = o.fill_bytes(o.FILL_PATTERN, 1, size, 7)
    b = o.fill_bytes(o.FILL_PATTERN, 2, size, 7)
    path = o.create_disk_layer(str(tmp_path), 0, 1, size, contents=a)
    xs = [o.Xfer(0, 1, 1, 0, 1000, size), o.Xfer(0, 1, 1, 1000, size - 1000, size), o.Xfer(0, 2, 2, 0, size, size)]
    secs, moved = co.tcp_run(3, 0, xs, [a, None, b], base_port=24100, src_paths=[path, path, None])
    assert moved == 2 * size
    with pytest.raises(RuntimeError):                      # the file does not hold what the caller expects: verification fails
        co.tcp_run(3, 0, [o.Xfer(0, 1, 1, 0, size, size)], [b], base_port=24200, src_paths=[path])
    rate = 20_000_000
    want_s = o.limiter_write_times(size, rate)[-1][0] / 1e9
    fast, _ = co.tcp_run(3, 0, xs[2:], [b], base_port=24300)
    paced, _ = co.tcp_run(3, 0, xs[2:], [b], base_port=24400, rates=[rate])
    assert want_s <= paced < want_s + 0.25 and fast < paced


def test_streaming_oracle_helpers_compose_the_two_definitions():
    """oracle_fill_checksum / oracle_fill_range (used on full-size layers by bench.py and the GPU tests) are the fill and the
    checksum composed; pinned here against the plain calls and the numpy restatement on ragged sizes."""
    import numpy as np
    from oracle import c_oracle as co
    for n in (0, 5, 262144, 262144 * 3 + 4321, 1_000_003):
        for kind in (o.FILL_ZERO, o.FILL_PATTERN, o.FILL_POISON):
            want = co.checksum(co.fill_bytes(kind, 7, n, 99))
            assert np.array_equal(co.fill_checksum(kind, 7, n, 99), want), (n, kind)
    n = 262144 * 3 + 4321
    assert np.array_equal(co.fill_checksum(o.FILL_PATTERN, 7, n, 99), o.checksum(o.fill_bytes(o.FILL_PATTERN, 7, n, 99)))
    whole = o.fill_bytes(o.FILL_PATTERN, 11, n, 5)
    for off, ln in ((0, 100), (8, 4096), (262144, 262144), (n - 4321 + 1 - 1 - (n - 4321) % 8, 77), (n - 8 * 3, 24)):
        off -= off % 8
        ln = min(ln, n - off)
        assert np.array_equal(co.fill_range(o.FILL_PATTERN, 11, off, ln, 5), whole[off:off + ln]), (off, ln)
    many = co.fill_checksum_many(o.FILL_PATTERN, [1, 2, 3], 300000, 4, threads=2)
    assert all(np.array_equal(many[l], o.checksum(o.fill_bytes(o.FILL_PATTERN, l, 300000, 4))) for l in (1, 2, 3))
