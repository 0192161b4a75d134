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


if __name__ == "__main__":
    main()

# plan_digests.json is produced from the C++ planners (the product), not from the oracle: it pins the plans of every
# BASELINE configuration x mode so that a later change to host/plan.cpp or host/flow.cpp that alters a plan is noticed.
# Regenerate with the snippet in tests/test_host_cpu.py::test_plan_digests_are_stable (set REGENERATE = True there).
