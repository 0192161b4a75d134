"""Differential tests of the plan generators: the C++ mirror (host/plan.cpp, host/flow.cpp, through dissem_plan_json)
against the independent numpy/Python restatement in oracle/, on RANDOM configurations (hypothesis, derandomised so the
suite is reproducible).  Either both reject a configuration or both produce the same transfers, t and self jobs."""
import json

import pytest
from hypothesis import HealthCheck, given, settings
from hypothesis import strategies as st

import dissem_b200 as d
from oracle import dissem_oracle as o


@st.composite
def configs(draw, need_bw=False):
    n_nodes = draw(st.integers(2, 6))
    n_layers = draw(st.integers(1, 7))
    sizes = [draw(st.sampled_from([1, 17, 4096, 262144, 262145, 1 << 20, 3_000_001, 67_108_864])) for _ in range(n_layers)]
    rates = st.sampled_from([1_000, 77_777, 1 << 20, 209_715_200, 900_000_000_000])
    nodes = []
    for i in range(n_nodes):
        held = [l for l in range(n_layers) if draw(st.integers(0, 9)) < (8 if i == 0 else 3)]
        nodes.append({"Id": i, "Addr": f"gpu:{i}", "NetworkBW": draw(rates), "IsLeader": i == 0,
                      "Sources": {"2": draw(rates) if need_bw or draw(st.booleans()) else 0},
                      "InitialLayers": {"2": {str(l): {"LayerSize": sizes[l]} for l in held}}})
    assignment = {}
    for i in range(n_nodes):
        want = [l for l in range(n_layers) if draw(st.integers(0, 9)) < 4]
        if want:
            assignment[str(i)] = {str(l): {} for l in want}
    return json.dumps({"Nodes": nodes, "Assignment": assignment})


def _rows(plan):
    return [tuple(x[:6]) for x in plan["xfers"]]


def _orows(xs):
    return [(x.src, x.dst, x.layer, x.off, x.size, x.total) for x in xs]


def _both(cpp, orc):
    """Run both sides; a rejection on one side must be a rejection on the other."""
    try:
        want = orc()
    except (ValueError, KeyError) as e:
        with pytest.raises(d.DissemError):
            cpp()
        return None, None
    return cpp(), want


COMMON = dict(max_examples=120, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow])


@settings(**COMMON)
@given(configs())
def test_mode0_random_configs(text):
    got, want = _both(lambda: d.plan(text, 0), lambda: o.plan_mode0(o.parse_config(text)))
    if got is not None:
        assert _rows(got) == _orows(want)


@settings(**COMMON)
@given(configs(), st.integers(0, 2**64 - 1), st.sampled_from([0, 1]))
def test_mode1_random_configs(text, seed, policy):
    got, want = _both(lambda: d.plan(text, 1, seed=seed, policy=policy),
                      lambda: o.plan_mode1(o.parse_config(text), seed=seed, policy="balanced" if policy else "random"))
    if got is not None:
        assert _rows(got) == _orows(want)
        cand = o.mode1_owner_candidates(o.parse_config(text))
        assert all(s in cand[(dst, l)] for (s, dst, l, *_r) in _rows(got))


@settings(**COMMON)
@given(configs())
def test_mode2_initial_jobs_random_configs(text):
    got, want = _both(lambda: d.plan(text, 2), lambda: o.plan_mode2_initial(o.parse_config(text)))
    if got is not None:
        assert _rows(got) == _orows(want)


@settings(**COMMON)
@given(configs(need_bw=True), st.sampled_from([1, 1000, 1_000_000]), st.sampled_from([0, 1, 2, 3, 6, 7]))
def test_mode3_random_configs(text, scale, flags):
    kw = dict(time_scale=scale, zero_rate_unlimited=bool(flags & 1), sizes_from_announces=bool(flags & 2), multi_dest=bool(flags & 4))
    got, want = _both(lambda: d.plan(text, 3, time_scale=scale, flags=flags), lambda: o.plan_mode3(o.parse_config(text), **kw))
    if got is not None:
        t, xs, self_jobs = want
        assert got["t"] == t and _rows(got) == _orows(xs)
        assert [tuple(j) for j in got["self_jobs"]] == [(n, l, s) for (n, l, s, _off) in self_jobs]
        # every fragment list tiles what it delivers, and rates are dataSize * ticks-per-second / t (node.go:1281)
        for x in got["xfers"]:
            assert x[6] == (x[4] * scale // t if t > 0 else 0)


def test_malformed_configs_are_rejected_not_crashed_on():
    """Mutated and truncated configurations: every call either plans or returns an error code; the library (which a Go
    process would have loaded) never aborts or hangs.  The reference ignores JSON errors (cmd/config.go:58-59)."""
    import random
    base = [o.config_readme(), o.config_ring_fixture(4, 5), o.config_conf_json(), o.config_stage_ring(4, 8, 1000)]
    rng = random.Random(1)
    junk = ['', '{', '}', '"', ':', ',', '[', ']', '-', '9' * 25, 'null', 'true', '1e999', '\\', '\\u12', '\x00']
    weird = [None, -1, 2**63, 2**64, 1.5, "x", [], {}, True, -2**63]

    def mutate(x):
        if isinstance(x, dict):
            for key in list(x):
                r = rng.random()
                if r < 0.05:
                    del x[key]
                elif r < 0.10:
                    x[key] = rng.choice(weird)
                else:
                    mutate(x[key])
        elif isinstance(x, list):
            for v in x:
                mutate(v)

    planned = rejected = 0
    for it in range(700):
        t = rng.choice(base)
        k = rng.randrange(4)
        if k == 0:
            i = rng.randrange(len(t))
            t = t[:i] + rng.choice(junk) + t[i + rng.randrange(3):]
        elif k == 1:
            t = t[:rng.randrange(len(t))]
        elif k == 2:
            obj = json.loads(t)
            mutate(obj)
            t = json.dumps(obj)
        else:
            t = ''.join(rng.choice('{}[]":,0123456789.eE-+ \n\tabcNodesId') for _ in range(rng.randrange(60)))
        for mode in (0, 1, 2, 3):
            try:
                d.plan(t, mode, seed=it, time_scale=rng.choice([1, 1000]))
                planned += 1
            except d.DissemError as e:
                assert e.code in (-1, -9, -10)
                rejected += 1
    assert planned > 200 and rejected > 200


@st.composite
def transfer_lists(draw):
    n_nodes = draw(st.integers(2, 8))
    xs = []
    for layer in range(draw(st.integers(1, 5))):
        total = draw(st.sampled_from([1, 4096, 262144, 300_001, 9 << 20, 70 << 20, 300 << 20]))
        src = draw(st.integers(0, n_nodes - 1))
        dsts = draw(st.lists(st.integers(0, n_nodes - 1).filter(lambda n: n != src), min_size=1, max_size=4, unique=True))
        whole = draw(st.booleans())
        off = 0 if whole else draw(st.integers(0, total - 1))
        size = total if whole else draw(st.integers(1, total - off))
        flags = d.XFER_FILE if draw(st.integers(0, 4)) == 0 else 0
        xs += [d.Xfer(src, dst, layer, flags, off, size, total) for dst in dsts]
    gpus = [draw(st.integers(0, 3)) for _ in range(n_nodes)]
    return xs, gpus


@settings(max_examples=150, deadline=None, derandomize=True, suppress_health_check=[HealthCheck.too_slow])
@given(transfer_lists(), st.integers(1, 9))
def test_relay_and_optimize_conserve_every_byte(case, pieces):
    """Whatever dissem_plan_relay / dissem_plan_optimize do to a plan, every destination still receives exactly the bytes
    it was to receive, once; and whoever forwards a range has received it in an earlier round (or is the original source)."""
    xs, gpus = case
    want = {}
    for x in xs:
        want.setdefault((x.dst_node, x.layer), []).append((x.off, x.off + x.size))
    for plan in (d.plan_relay(xs, pieces), d.plan_optimize(xs, gpus)[0]):
        got = {}
        for y in plan:
            got.setdefault((y.dst_node, y.layer), []).append((y.off, y.off + y.size, d.xfer_round(y), y.src_node, y.flags))
        assert set(got) == set(want)
        for key, ranges in got.items():
            ranges.sort()
            (a0, b0), = want[key]
            assert ranges[0][0] == a0 and ranges[-1][1] == b0
            assert all(r1[1] == r2[0] for r1, r2 in zip(ranges[:-1], ranges[1:]))      # tiles, no gap, no overlap
        orig_src = {x.layer: x.src_node for x in xs}
        for y in plan:
            if y.src_node == orig_src[y.layer]:
                continue
            # a forwarder: it must itself receive [off, off+size) of this layer in an earlier round
            feeds = [r for r in got[(y.src_node, y.layer)] if r[0] <= y.off and y.off + y.size <= r[1]]
            assert feeds and all(r[2] < d.xfer_round(y) or (r[4] & d.XFER_FILE) for r in feeds)
            assert not y.flags & d.XFER_FILE
