"""CPU tests of the product's host side: the C ABI loads and exports every declared symbol, the C++ plan
generators agree with the oracle on every config/mode, and the control-plane wire format matches the
reference's.  No compute entry point is called: there is no GPU here."""
import importlib
import json
import os
import re

import pytest

import dissem_b200 as d
from oracle import dissem_oracle as o

host = importlib.import_module("distributed-llm-dissemination_b200.host")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = []
    for hdr in ("dissem.h", "dissem_host.h"):
        text = open(os.path.join(ROOT, "include", hdr)).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names += re.findall(r"\b(dissem_[a-z0-9_]+)\s*\(", text)
    return sorted(set(names))


def test_library_loads_and_exports_every_declared_symbol():
    L = d.load_library()
    syms = _declared_symbols()
    assert len(syms) >= 30
    for name in syms:
        assert hasattr(L, name), f"{name} declared in include/ but not exported"
    assert L.dissem_version() == 1


def test_no_cpu_fallback_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert d.lib().dissem_device_count() == 0
    with pytest.raises(d.DissemError) as e:
        d.Engine([0, 0])
    assert e.value.code == -2 and "no CPU path" in str(e.value)
    with pytest.raises(d.DissemError):
        host.run_distribution(o.config_readme(), 0)


def _rows(plan):
    return [tuple(x[:6]) for x in plan["xfers"]]


def _orows(xs):
    return [(x.src, x.dst, x.layer, x.off, x.size, x.total) for x in xs]


CONFIGS = {
    "readme": o.config_readme,
    "identity": o.config_identity_fixture,
    "ring": o.config_ring_fixture,
    "stage_ring_8x32": lambda: o.config_stage_ring(8, 32, o.S2),
    "stage_seed_8x32_8B": lambda: o.config_stage_ring(8, 32, o.S8B, ring=False),
    "stage_ring_4x8": lambda: o.config_stage_ring(4, 8, 1 << 20),
    "stage_ring_2x8": lambda: o.config_stage_ring(2, 8, 12345),
}


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_mode0_plan_matches_oracle(name):
    text = CONFIGS[name]()
    assert _rows(d.plan(text, 0)) == _orows(o.plan_mode0(o.parse_config(text)))


@pytest.mark.parametrize("name", sorted(CONFIGS))
@pytest.mark.parametrize("policy", [0, 1])
def test_mode1_plan_matches_oracle(name, policy):
    text = CONFIGS[name]()
    for seed in (0, 1, 0x5EED, 2**63 + 11):
        got = _rows(d.plan(text, 1, seed=seed, policy=policy))
        want = _orows(o.plan_mode1(o.parse_config(text), seed=seed, policy="balanced" if policy else "random"))
        assert got == want
        cand = o.mode1_owner_candidates(o.parse_config(text))      # and every pick is a valid owner
        for (s, dst, l, *_rest) in got:
            assert s in cand[(dst, l)]


def test_mode1_fanout_config():
    text = o.config_fanout()
    got = _rows(d.plan(text, 1, seed=3))
    assert got == _orows(o.plan_mode1(o.parse_config(text), seed=3)) and len(got) == 239
    with pytest.raises(d.DissemError) as e:
        d.plan(text, 0)
    assert e.value.code == -10


def test_mode3_flow_kat_and_oracle_agreement():
    p = d.plan(o.config_conf_json(), 3)
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "flow_conf_json.json")))
    assert p["t"] == gold["t"] == 60
    assert [list(r) for r in _rows(p)] == gold["xfers"]
    assert all(x[6] == x[4] // 60 for x in p["xfers"])          # rate = dataSize / minTime, node.go:1281


@pytest.mark.parametrize("scale", [1, 1000, 1_000_000])
@pytest.mark.parametrize("name", ["stage_seed_8x32_8B", "stage_ring_8x32", "stage_ring_4x8"])
def test_mode3_matches_oracle_across_time_scales(name, scale):
    text = CONFIGS[name]()
    p = d.plan(text, 3, time_scale=scale)
    t, xs, self_jobs = o.plan_mode3(o.parse_config(text), time_scale=scale)
    assert p["t"] == t and _rows(p) == _orows(xs)
    assert [tuple(j) for j in p["self_jobs"]] == [(n, l, s) for (n, l, s, _off) in self_jobs]


def test_mode3_zero_rate_and_sizes_extensions():
    text = o.config_readme()                                    # no Sources: LimitRate 0 everywhere
    with pytest.raises(d.DissemError):
        d.plan(text, 3)                                         # "tUpper not found", flow.go:162-165
    cfg = json.loads(text)
    for n in cfg["Nodes"]:
        n["NetworkBW"] = 1000
    text = json.dumps(cfg)
    p = d.plan(text, 3, flags=1)
    t, xs, _ = o.plan_mode3(o.parse_config(text), zero_rate_unlimited=True)
    assert p["t"] == t and _rows(p) == _orows(xs)


def test_bad_inputs_fail_loudly():
    with pytest.raises(d.DissemError) as e:
        d.plan("{not json", 0)
    assert e.value.code == -9
    with pytest.raises(d.DissemError) as e:
        d.plan(json.dumps({"Nodes": [{"Id": 0}], "Assignment": {}}), 0)
    assert e.value.code == -10 and "no leader" in str(e.value)
    with pytest.raises(d.DissemError):
        d.plan(o.config_readme(), 7)                            # "unknown mode", cmd/main.go:163-164


def test_schema_detection():
    assert d.plan(o.config_readme(), 0)["flat_schema"] is True
    assert d.plan(o.config_conf_json(), 3)["flat_schema"] is False
    # Go's decoder is case-insensitive on field names: "Id" fills ID (cmd/config.go:22)
    text = o.config_readme().replace('"Id"', '"ID"').replace('"IsLeader"', '"isleader"')
    assert _rows(d.plan(text, 0)) == _rows(d.plan(o.config_readme(), 0))


# ---- wire format, SURVEY Appendix A ----------------------------------------------------------------
def test_wire_format_known_answers():
    w, dec = host.msg_roundtrip(kind="announce", SrcID=1,
                                LayerIDs={"4": {"Location": 0, "LimitRate": 0, "SourceType": 0}})
    assert w == '{"type":0,"src":"1","payload":{"SrcID":1,"LayerIDs":{"4":{"Location":0,"LimitRate":0,"SourceType":0}}}}'
    w, dec = host.msg_roundtrip(kind="ack", SrcID=1, LayerID=1, location=1)
    assert w == '{"type":1,"src":"1","payload":{"SrcID":1,"LayerID":1}}'
    assert dec["location_after_decode"] == 0          # unexported field: always decodes as in-memory (message.go:62-66)
    w, _ = host.msg_roundtrip(kind="retransmit", SrcID=0, LayerID=1, DestID=2)
    assert w == '{"type":3,"src":"0","payload":{"SrcID":0,"LayerID":1,"DestID":2}}'
    w, _ = host.msg_roundtrip(kind="flowRetransmit", SrcID=0, LayerID=1, DestID=7, DataSize=5, Offset=6, Rate=7)
    assert w == '{"type":4,"src":"0","payload":{"SrcID":0,"LayerID":1,"DestID":7,"DataSize":5,"Offset":6,"Rate":7}}'
    w, _ = host.msg_roundtrip(kind="startup", SrcID=0)
    assert w == '{"type":6,"src":"0","payload":{"SrcID":0}}'
    w, dec = host.msg_roundtrip(kind="simple", SrcAddr="a:1", PayloadStr='he said "hi"\n')
    assert json.loads(w)["payload"] == {"SrcAddr": "a:1", "PayloadStr": 'he said "hi"\n'} and dec["type"] == 7
    w, dec = host.msg_roundtrip(kind="layerHeader", SrcID=0, LayerID=1, LayerSize=10, TotalSize=30, Offert=20)
    assert w == '{"type":2,"src":"0","payload":{"SrcID":0,"LayerID":1,"LayerSize":10,"TotalSize":30,"Offert":20}}'
    assert dec["payload"]["Offert"] == 20             # the offset survives (the reference drops it, transport.go:221)


def test_plan_directions_heuristic():
    """dissem_plan_directions: pull only where the reverse NVLink ports are (nearly) idle."""
    text = o.config_stage_ring(8, 32, o.S2)
    xs = d.parse_plan(d.plan(text, 1, seed=0x5EED))
    one = d.plan_directions(xs, [0] * 8)
    assert all(x.flags == 0 for x in one)                              # one GPU: nothing crosses a link
    two = d.plan_directions(xs, [n % 2 for n in range(8)])
    assert all(x.flags == 0 for x in two)                              # both directions loaded: push
    eight = d.plan_directions(xs, list(range(8)))
    pulled = [(x.src_node, x.dst_node) for x in eight if x.flags & d.XFER_PULL]
    assert pulled and all(s == 0 for s, _ in pulled)                   # the leader's egress dominates, its ingress is idle
    assert all(x.flags == 0 for x in eight if x.src_node != 0)
    # a single seed GPU: every transfer is pulled
    seed = d.parse_plan(d.plan(o.config_stage_ring(8, 32, o.S2, ring=False), 0))
    assert all(x.flags & d.XFER_PULL for x in d.plan_directions(seed, list(range(8))))
    # idempotent, and the flag is cleared when the mapping changes
    again = d.plan_directions(eight, [0] * 8)
    assert all(x.flags == 0 for x in again)
    with pytest.raises(d.DissemError):
        d.plan_directions(xs, [0, 1])


def test_plan_relay_structure():
    """dissem_plan_relay: every destination still receives every byte exactly once, the source sends each byte once,
    and a hop forwards a piece exactly one round after it arrived."""
    S = 10 * d.CHUNK_BYTES + 100
    xs = [d.Xfer(0, 1, 5, 0, 0, S, S), d.Xfer(0, 2, 5, 0, 0, S, S), d.Xfer(0, 3, 5, 0, 0, S, S),
          d.Xfer(2, 3, 9, 0, 0, S, S), d.Xfer(1, 0, 7, 0, 4096, 3 * d.CHUNK_BYTES, S)]
    r = d.plan_relay(xs, 4)
    cover = {}
    for x in r:
        assert x.size > 0 and x.off % d.CHUNK_BYTES == 0 or x.off == 4096
        cover.setdefault((x.dst_node, x.layer), []).append((x.off, x.size, x.src_node, d.xfer_round(x)))
    for (dst, layer), want in {(1, 5): (0, S), (2, 5): (0, S), (3, 5): (0, S), (3, 9): (0, S), (0, 7): (4096, 3 * d.CHUNK_BYTES)}.items():
        pos = want[0]
        for off, size, _src, _rnd in sorted(cover[(dst, layer)]):
            assert off == pos
            pos += size
        assert pos == want[0] + want[1]
    sent_by_source = sum(x.size for x in r if x.layer == 5 and x.src_node == 0)
    assert sent_by_source == S                                             # once, not three times
    # chain: the piece a node forwards in round r arrived at it in round r-1
    arrived = {(x.dst_node, x.layer, x.off): d.xfer_round(x) for x in r}
    for x in r:
        if x.layer == 5 and x.src_node != 0:
            assert arrived[(x.src_node, 5, x.off)] == d.xfer_round(x) - 1
    assert max(d.xfer_round(x) for x in r) == 4 + 3 - 2                    # pieces + hops - 2 (rounds count from 0)
    assert all(x.flags & d.XFER_PULL == 0 for x in r)
    with pytest.raises(d.DissemError):
        d.plan_relay(xs, 0)


@pytest.mark.parametrize("name", sorted(CONFIGS))
def test_mode2_initial_jobs_match_oracle(name):
    """PullRetransmitLeaderNode's initial job table (node.go:810-886): rarest layer first, least-loaded owner."""
    text = CONFIGS[name]()
    got = _rows(d.plan(text, 2))
    assert got == _orows(o.plan_mode2_initial(o.parse_config(text)))
    cand = o.mode1_owner_candidates(o.parse_config(text))
    assert len(got) == len(cand)
    for (s, dst, l, *_r) in got:
        assert s in cand[(dst, l)]


def test_mode2_ring_fixture_known_answer():
    """node_test.go:246-272 (tcp_pullretransmission) on the ring placement: every needed layer has owners {leader, one
    peer}; all loads start at 0, so ties go to the lowest id (node.go:967-971): layer 1 -> leader, then the peers."""
    got = _rows(d.plan(o.config_ring_fixture(), 2))
    assert got == [(0, 1, 1, 0, 1, 1), (3, 2, 2, 0, 1, 1), (4, 3, 3, 0, 1, 1), (1, 4, 4, 0, 1, 1)]


def test_mode3_multi_destination_extension():
    """Per-(layer, destination) demand (the reference allows one destination per layer, node.go:1078): every
    destination's fragments tile its copy of the layer; C++ == oracle; single-destination plans are unchanged."""
    text = o.config_fanout(n_nodes=8, n_layers=24, size=3 * (1 << 20) + 4096)
    flags = host.PLAN_SIZES_FROM_ANNOUNCES | host.PLAN_MULTI_DEST
    p = d.plan(text, 3, time_scale=1_000_000, flags=flags)
    cfg = o.parse_config(text)
    t, xs, _self = o.plan_mode3(cfg, time_scale=1_000_000, sizes_from_announces=True, multi_dest=True)
    assert p["t"] == t and _rows(p) == _orows(xs)
    need = {(n, l) for n, ls in cfg.assignment.items() for l in ls}
    got = {}
    for (s, dst, l, off, size, total) in _rows(p):
        got.setdefault((dst, l), []).append((off, size))
        assert l in cfg.node(s).layers
    assert set(got) == need
    for key, frs in got.items():
        pos = 0
        for off, size in sorted(frs):
            assert off == pos
            pos += size
        assert pos == 3 * (1 << 20) + 4096
    # one destination per layer: the extension changes nothing
    text1 = o.config_stage_ring(8, 32, o.S2)
    assert _rows(d.plan(text1, 3, time_scale=1_000_000, flags=host.PLAN_MULTI_DEST)) == _rows(d.plan(text1, 3, time_scale=1_000_000))
    assert _rows(d.plan(o.config_conf_json(), 3, flags=host.PLAN_MULTI_DEST)) == _rows(d.plan(o.config_conf_json(), 3))


def test_plan_optimize_picks_relay_only_when_it_unloads_the_busiest_port():
    size = 1_711_308_800

    def fanout(one_seed, n=8, layers=80):
        xs = []
        for l in range(layers):
            seed = 0 if one_seed else (l + 4) % n
            dsts = {l % n, (l + 1) % n} | ({(l + 2) % n} if l % 3 >= 1 else set()) | ({(l + 3) % n} if l % 3 == 2 else set())
            xs += [d.Xfer(seed, dst, l, 0, 0, size, size) for dst in sorted(dsts - {seed})]
        return xs

    seed_plan, pieces = d.plan_optimize(fanout(True), list(range(8)))
    assert pieces == 25 and max(d.xfer_round(x) for x in seed_plan) > 0
    assert sum(x.size for x in seed_plan if x.src_node == 0) == 80 * size          # the seed sends every layer once
    assert all(x.flags & d.XFER_PULL for x in seed_plan if x.src_node == 0)        # and its hops are pulled
    bal_plan, pieces = d.plan_optimize(fanout(False), list(range(8)))
    assert pieces == 0 and len(bal_plan) == 239 and all(d.xfer_round(x) == 0 for x in bal_plan)
    # everything on one GPU: nothing to optimise
    one, pieces = d.plan_optimize(fanout(True), [0] * 8)
    assert pieces == 0 and all(x.flags == 0 for x in one)


def test_plan_align_moves_cuts_to_chunk_boundaries_and_keeps_coverage():
    """dissem_plan_align on flow plans (mode 3 cuts a layer wherever the max flow falls, flow.go:193-212): afterwards no two
    fragments share a 256 KiB checksum chunk, and every (destination, layer) is still covered exactly once."""
    C = d.CHUNK_BYTES
    for text, scale in ((o.config_stage_ring(8, 32, o.S2), 1_000_000), (o.config_stage_ring(8, 32, o.S2, ring=False), 1_000_000),
                        (o.config_conf_json(), 1), (o.config_stage_ring(4, 8, (1 << 20) + 52), 1_000_000)):
        xs = d.parse_plan(d.plan(text, 3, time_scale=scale))
        al = d.plan_align(xs)
        cover_before, cover_after = {}, {}
        for lst, cov in ((xs, cover_before), (al, cover_after)):
            for x in lst:
                cov.setdefault((x.dst_node, x.layer, x.total), []).append((x.off, x.off + x.size))
        assert set(cover_before) == set(cover_after)
        for key, spans in cover_after.items():
            spans.sort()
            assert spans[0][0] == min(a for a, _b in cover_before[key]) and spans[-1][1] == max(b for _a, b in cover_before[key])
            for (a0, a1), (b0, b1) in zip(spans, spans[1:]):
                assert a1 == b0 and b0 % C == 0                              # contiguous, cut on a chunk boundary
        assert all(x.size > 0 for x in al)
        # every fragment still comes from a node that holds the layer (any sender holds the WHOLE layer)
        holders = {(x.src_node, x.layer) for x in xs}
        assert all((x.src_node, x.layer) in holders for x in al)
    # a fragment smaller than the move disappears; unrelated and non-contiguous transfers are left alone
    tiny = [d.Xfer(0, 3, 7, 0, 0, C + 100, 4 * C), d.Xfer(1, 3, 7, 0, C + 100, 50, 4 * C), d.Xfer(2, 3, 7, 0, C + 150, 3 * C - 150, 4 * C),
            d.Xfer(0, 4, 7, 0, 10, 20, 4 * C), d.Xfer(1, 4, 7, 0, 100, 20, 4 * C)]
    al = d.plan_align(tiny)
    rows = sorted((x.src_node, x.dst_node, x.off, x.size) for x in al)
    assert rows == sorted([(0, 3, 0, C), (2, 3, C, 3 * C), (0, 4, 10, 20), (1, 4, 100, 20)])


def test_plan_localize_takes_same_gpu_copies_and_reaches_the_best_sender_port_loads():
    """BASELINE config 2 (ring placement) with its 8 nodes on 2 / 4 / 8 GPUs: after dissem_plan_localize + dissem_plan_optimize
    the busiest NVLink port carries what the best sender choice allows (805 / 537 / 268 MB), and the
    delivered store is the one the untouched flow plan delivers (oracle.apply_transfers on small layers)."""
    text = o.config_stage_ring(8, 32, o.S2, ring=True)
    p = d.plan(text, 3, time_scale=1_000_000)
    xs = d.parse_plan(p)
    holders = [(nd["id"], int(l)) for nd in p["nodes"] for l in nd["layers"]]
    want = {2: 3 * 4 * o.S2, 4: 2 * 4 * o.S2, 8: 4 * o.S2}
    for n_gpus, best in want.items():
        g = [n % n_gpus for n in range(8)]
        xo, _ = d.plan_optimize(d.plan_localize(xs, g, holders), g)
        eg, ing = [0] * n_gpus, [0] * n_gpus
        for x in xo:
            assert (x.src_node, x.layer) in set(holders)
            if g[x.src_node] != g[x.dst_node]:
                eg[g[x.src_node]] += x.size
                ing[g[x.dst_node]] += x.size
        assert max(eg + ing) <= best + 2 * d.CHUNK_BYTES, (n_gpus, eg, ing)
        assert sum(x.size for x in xo) == sum(x.size for x in xs)
    # same store as the plain plan, byte for byte (small ragged layers so that the numpy oracle finishes at once)
    small = o.config_stage_ring(8, 32, 3 * d.CHUNK_BYTES + 1234, ring=True)
    ps = d.plan(small, 3, time_scale=1_000_000)
    plain = d.parse_plan(ps)
    hs = [(nd["id"], int(l)) for nd in ps["nodes"] for l in nd["layers"]]
    g = [n % 2 for n in range(8)]
    tuned, _ = d.plan_optimize(d.plan_localize(plain, g, hs), g)

    def store_after(xfers):
        st = {(n, l): o.fill_bytes(o.FILL_PATTERN, l, 3 * d.CHUNK_BYTES + 1234, 5) for (n, l) in hs}
        o.apply_transfers(st, [o.Xfer(x.src_node, x.dst_node, x.layer, x.off, x.size, x.total) for x in xfers])
        return st

    a, b = store_after(plain), store_after(tuned)
    assert set(a) == set(b)
    for k in a:
        assert (a[k] == b[k]).all(), k


def test_plan_digests_are_stable():
    """Every BASELINE configuration x mode: the C++ planners still produce the committed plans (tests/golden/plan_digests.json),
    and the oracle agrees with each of them where it restates the mode."""
    import hashlib
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "plan_digests.json")))
    cases = {
        "readme": (o.config_readme(), 1), "identity": (o.config_identity_fixture(), 0), "ring": (o.config_ring_fixture(), 0),
        "config2_ring_8x32x64MiB": (o.config_stage_ring(8, 32, o.S2), 0),
        "config3_llama8b": (o.config_stage_ring(8, 32, o.S8B, ring=False), 0),
        "config4_fanout_70b": (o.config_fanout(), 2 | 4),
        "config5_seed_8x32x64MiB": (o.config_stage_ring(8, 32, o.S2, ring=False), 0),
        "conf_json": (o.config_conf_json(), 0),
    }
    seen = 0
    for name, (text, flags) in cases.items():
        for mode in (0, 1, 2, 3):
            for ts in ((1, 1_000_000) if mode == 3 else (1,)):
                key = f"{name}/m{mode}" + (f"/ts{ts}" if mode == 3 else "")
                want = gold[key]
                try:
                    p = d.plan(text, mode, seed=0x5EED, policy=0, time_scale=ts, flags=flags)
                except d.DissemError as e:
                    assert want == {"error": e.code}, key
                    continue
                rows = [x[:6] for x in p["xfers"]]
                got = {"n": len(rows), "bytes": sum(r[4] for r in rows), "t": p["t"],
                       "sha256": hashlib.sha256(json.dumps(rows).encode()).hexdigest()}
                assert got == want, key
                seen += 1
    assert seen >= 30 and len(gold) == 40


@pytest.mark.parametrize("off,size,rate", [
    (0, 1, 1000), (0, 262144, 1 << 20), (0, 262145, 1 << 20), (0, 5 * 262144 + 100, 20_000_000),
    (13, 3 * 262144, 7_000_000), (262144 - 1, 2, 1 << 20), (5 * 262144 + 77, 40 * 262144 + 5, 3_000_000_000),
    (100_003, 64 << 20, 200_000_000_000), (0, 1 << 30, 900_000_000_000),
])
def test_pace_schedule_never_runs_ahead_of_the_reference_limiter(off, size, rate):
    """writeWithLimit (transport.go:407-424) over x/time/rate's token bucket, restated in the oracle: a piece is released
    no earlier than the sub-write that carries its last byte, and the last piece exactly when the last sub-write is."""
    pieces = host.pace_schedule(off, size, rate)
    writes = o.limiter_write_times(size, rate)
    assert pieces[0][0] == off and sum(p[1] for p in pieces) == size
    for (a, n, _), (b, _, _) in zip(pieces[:-1], pieces[1:]):
        assert a + n == b and b % 262144 == 0                   # pieces tile the fragment and end on checksum chunks
    tol = len(writes) + 2                                       # the limiter truncates every wait to whole ns
    starts = [w[1] for w in writes]
    import bisect
    for (a, n, t_ns) in pieces:
        last_byte = a + n - 1 - off
        w = writes[bisect.bisect_right(starts, last_byte) - 1]
        assert w[1] <= last_byte < w[1] + w[2]
        assert abs(t_ns - w[0]) <= tol, (a, n, t_ns, w)
    assert abs(pieces[-1][2] - writes[-1][0]) <= tol
    assert all(x[2] <= y[2] for x, y in zip(pieces[:-1], pieces[1:]))
    # launches are spaced: at most ~4000 a second unless the reference's own 256 KiB quantum is coarser
    if len(pieces) > 2:
        gaps = [y[2] - x[2] for x, y in zip(pieces[1:-2], pieces[2:-1])]
        assert all(g >= min(240_000, int(262144 / rate * 1e9) - 1) for g in gaps)


def test_pace_schedule_edge_cases():
    assert host.pace_schedule(0, 0, 1000) == []
    assert host.pace_schedule(7, 1 << 20, 0) == [(7, 1 << 20, 0)]          # 0 = unlimited here (the reference would hang)
    assert host.pace_schedule(7, 1 << 20, -5) == [(7, 1 << 20, 0)]
    with pytest.raises(host.HostError):
        host.pace_schedule(-1, 10, 10)
    with pytest.raises(ValueError):
        o.limiter_write_times(10, 0)


def test_disk_sources_are_never_pulled_or_relayed():
    """DISSEM_XFER_FILE: the node that holds the file streams it; the scheduling decisions leave such transfers alone."""
    multi = importlib.import_module("distributed-llm-dissemination_b200.multi")
    S = 8 << 20
    xs = [d.Xfer(0, n, 7, d.XFER_FILE, 0, S, S) for n in (1, 2, 3)] + [d.Xfer(0, n, 8, 0, 0, S, S) for n in (1, 2, 3)]
    out = d.plan_directions(xs, [0, 1, 2, 3])
    assert [x.flags for x in out[:3]] == [d.XFER_FILE] * 3               # single seed: everything else is pulled
    assert all(x.flags & d.XFER_PULL for x in out[3:])
    relayed = d.plan_relay(xs, 4)
    files = [x for x in relayed if x.flags & d.XFER_FILE]
    assert [(x.src_node, x.dst_node, x.off, x.size, x.flags) for x in files] == [(0, n, 0, S, d.XFER_FILE) for n in (1, 2, 3)]
    assert any(x.src_node != 0 for x in relayed if x.layer == 8)         # the HBM-resident layer is chained
    owner = multi.owner_round_robin(4)
    both = d.Xfer(0, 1, 7, d.XFER_FILE | d.XFER_PULL, 0, S, S)
    assert multi.my_transfers([both], owner, 0) == [both] and multi.my_transfers([both], owner, 1) == []
    assert multi.imports_needed([both], owner, 0) == [(1, 7, S)] and multi.exports_needed([both], owner, 1) == [(1, 7, S)]


def test_plan_optimize_reads_a_disk_layer_once():
    """A layer on disk that several nodes need: one transfer keeps DISSEM_XFER_FILE, the other destinations are fed from
    the first one's HBM in later rounds; every destination still receives every byte exactly once."""
    S = 200 << 20
    xs = [d.Xfer(0, n, 7, d.XFER_FILE, 0, S, S) for n in (1, 2, 3)] + [d.Xfer(0, 0, 9, d.XFER_FILE, 0, S, S), d.Xfer(0, 2, 9, d.XFER_FILE, 0, S, S)]
    for gpus in ([0, 1, 2, 3], [0, 0, 0, 0]):
        out, _ = d.plan_optimize(xs, gpus)
        files = [(x.src_node, x.dst_node, x.layer) for x in out if x.flags & d.XFER_FILE]
        assert files == [(0, 1, 7), (0, 0, 9), (0, 2, 9)]                 # a node loading its own copy feeds nobody
        assert all(d.xfer_round(x) == 0 for x in out if x.flags & d.XFER_FILE)
        rest = [x for x in out if not x.flags & d.XFER_FILE]
        assert rest and all(d.xfer_round(x) >= 1 and x.src_node in (1, 2, 3) and x.layer == 7 for x in rest)
        for dst in (2, 3):
            got = sorted((x.off, x.size) for x in rest if x.dst_node == dst)
            assert got[0][0] == 0 and sum(sz for _, sz in got) == S
            assert all(a + n == b for (a, n), (b, _) in zip(got[:-1], got[1:]))
    # without disk sources nothing is shifted
    plain, _ = d.plan_optimize([d.Xfer(0, n, 7, 0, 0, 1 << 20, 1 << 20) for n in (1, 2)], [0, 1, 2])
    assert all(d.xfer_round(x) == 0 for x in plain)


# ---- the reference's own transport tests (distributor/transport_test.go), replayed on the device transport ----------
def _simple(text, src="gpu:1"):
    return {"kind": "simple", "SrcAddr": src, "PayloadStr": text}


def test_transport_send_single():
    """TestTransportSendSingle (transport_test.go:18-68): p1.Send(peer2, msg) is delivered on p2.Deliver()."""
    r = host.transport_script([1, 2], [{"op": "send", "from": 1, "to": 2, "msg": _simple("hi from peer1")}])
    assert r["results"] == [0]
    assert [m["payload"]["PayloadStr"] for m in r["delivered"]["2"]] == ["hi from peer1"]
    assert r["delivered"]["1"] == [] and r["addresses"] == {"1": "gpu:1", "2": "gpu:2"}


def test_transport_send_three_in_order():
    """TestInmemoryTransportSendThree (transport_test.go:70-138): three messages arrive, equal to what was sent, in order."""
    texts = ["hi 0", "hi 1", "hi 2"]
    r = host.transport_script([1, 2], [{"op": "send", "from": 1, "to": 2, "msg": _simple(t)} for t in texts])
    assert r["results"] == [0, 0, 0]
    got = r["delivered"]["2"]
    assert [m["payload"]["PayloadStr"] for m in got] == texts
    assert all(m["type"] == 7 and m["src"] == "gpu:1" and m["payload"]["SrcAddr"] == "gpu:1" for m in got)
    # interleaved senders keep their own order (one queue per receiver)
    ops = [{"op": "send", "from": 1 + (i % 2), "to": 3, "msg": _simple(f"m{i}", src=f"gpu:{1 + i % 2}")} for i in range(10)]
    got = host.transport_script([1, 2, 3], ops)["delivered"]["3"]
    assert [m["payload"]["PayloadStr"] for m in got] == [f"m{i}" for i in range(10)]


def test_transport_broadcast_single():
    """TestInmemoryTransportBroadcastSingle (transport_test.go:140-186): Broadcast reaches the other peer — and, as
    transport.go:290-306 iterates the whole registry, the sender itself."""
    r = host.transport_script([1, 2, 3], [{"op": "broadcast", "from": 1, "msg": _simple("broadcast value")}])
    assert r["results"] == [0]
    for peer in ("1", "2", "3"):
        assert [m["payload"]["PayloadStr"] for m in r["delivered"][peer]] == ["broadcast value"]


def test_transport_contract_edges():
    """Self-send is legal; an unknown destination is an error the caller logs (transport.go:263-265);
    a closed peer is gone from the registry; RegisterPipe is outside the device transport; control messages keep what
    the JSON envelope keeps (ackMsg.location is unexported and therefore lost, message.go:62-66)."""
    ack = {"kind": "ack", "SrcID": 2, "LayerID": 5, "location": 1}
    r = host.transport_script([1, 2], [
        {"op": "send", "from": 1, "to": 1, "msg": _simple("to myself")},
        {"op": "send", "from": 1, "to": 9, "msg": _simple("nobody")},
        {"op": "send", "from": 2, "to": 1, "msg": ack},
        {"op": "pipe", "from": 1, "layer": 3, "to": 2},
        {"op": "close", "peer": 2},
        {"op": "send", "from": 1, "to": 2, "msg": _simple("too late")},
    ])
    assert r["results"] == [0, -4, 0, -8, 0, -4]
    one = r["delivered"]["1"]
    assert one[0]["payload"]["PayloadStr"] == "to myself"
    assert one[1]["type"] == 1 and one[1]["payload"] == {"SrcID": 2, "LayerID": 5} and one[1]["location_after_decode"] == 0
    assert r["delivered"]["2"] == []
    with pytest.raises(host.HostError):
        host.transport_script([1], [{"op": "send", "from": 1, "to": 1, "msg": {"kind": "nonsense"}}])
