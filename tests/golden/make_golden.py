"""Regenerates the committed golden fixtures from the CPU oracle:
    python tests/golden/make_golden.py
The fixtures let the GPU box (which has no /root/reference) and later rounds detect any drift
in the oracle's definitions.  checksum_kat.json: fill + checksum known answers.
flow_conf_json.json: the mode-3 plan for conf/config.json (t = 60)."""
import json
import os
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import dissem_oracle as o  # noqa: E402


def main():
    cases = []
    for kind, layer, size, seed in [(1, 0, 16, 0), (1, 1, 1, 0x5EED), (1, 3, 4099, 0x5EED), (0, 2, 262145, 0),
                                    (2, 2, 600_000, 0), (1, 31, 786_432 + 13, 0x5EED), (1, 79, 262_144, 1)]:
        data = o.fill_bytes(kind, layer, size, seed)
        cs = o.checksum(data)
        cases.append({"kind": kind, "layer": layer, "size": size, "seed": seed,
                      "head_hex": data[:16].tobytes().hex(),
                      "csum": [[int(a), int(b)] for a, b in cs.tolist()], "digest": o.digest(cs)})
    json.dump({"chunk_bytes": o.CHUNK_BYTES, "cases": cases},
              open(os.path.join(HERE, "checksum_kat.json"), "w"), indent=1)
    cfg = o.parse_config(o.config_conf_json())
    t, xs, _ = o.plan_mode3(cfg)
    json.dump({"t": t, "xfers": [[x.src, x.dst, x.layer, x.off, x.size, x.total] for x in xs]},
              open(os.path.join(HERE, "flow_conf_json.json"), "w"), indent=1)


def plan_digests():
    """plan_digests.json comes from the C++ planners (the product), not from the oracle: it pins the plan of every BASELINE
    configuration x mode, so that a later change to host/plan.cpp or host/flow.cpp that alters a plan is noticed
    (tests/test_host_cpu.py::test_plan_digests_are_stable).  Needs libdissem.so built; no GPU."""
    import hashlib

    import dissem_b200 as d
    cases = {
        "readme": (o.config_readme(), 1), "identity": (o.config_identity_fixture(), 0), "ring": (o.config_ring_fixture(), 0),
        "config2_ring_8x32x64MiB": (o.config_stage_ring(8, 32, o.S2), 0),
        "config3_llama8b": (o.config_stage_ring(8, 32, o.S8B, ring=False), 0),
        "config4_fanout_70b": (o.config_fanout(), 2 | 4),
        "config5_seed_8x32x64MiB": (o.config_stage_ring(8, 32, o.S2, ring=False), 0),
        "conf_json": (o.config_conf_json(), 0),
    }
    out = {}
    for name, (text, flags) in cases.items():
        for mode in (0, 1, 2, 3):
            for ts in ((1, 1_000_000) if mode == 3 else (1,)):
                key = f"{name}/m{mode}" + (f"/ts{ts}" if mode == 3 else "")
                try:
                    p = d.plan(text, mode, seed=0x5EED, policy=0, time_scale=ts, flags=flags)
                    rows = [x[:6] for x in p["xfers"]]
                    out[key] = {"n": len(rows), "bytes": sum(r[4] for r in rows), "t": p["t"],
                                "sha256": hashlib.sha256(json.dumps(rows).encode()).hexdigest()}
                except d.DissemError as e:
                    out[key] = {"error": e.code}
    json.dump(out, open(os.path.join(HERE, "plan_digests.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
    plan_digests()
