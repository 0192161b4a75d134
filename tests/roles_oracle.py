"""Helper for the role-driven GPU tests: run a distribution through the C++ roles with "dump_tables" and hold what
every assigned (node, layer) ended with — fused checksum table, received counter, first / last 4 KiB — against the CPU
oracle's store (oracle.apply_transfers over the transfers the run reports), instead of trusting the report's own
`verified` flag, which only compares values libdissem.so computed with each other."""
import importlib

import numpy as np

from oracle import dissem_oracle as o

host = importlib.import_module("distributed-llm-dissemination_b200.host")


def check_against_oracle(text: str, r: dict, seed: int = 0, fill: int = o.FILL_PATTERN, overrides=None) -> int:
    """Returns the number of received layers checked.  `r` must come from a run with dump_tables=True.
    overrides: {(node, layer): bytes} for initial copies that do not hold the seeded fill (a pre-existing layer file)."""
    cfg = o.parse_config(text)
    store = o.initial_store(cfg, fill, seed)
    for key, data in (overrides or {}).items():
        store[key] = np.array(data, dtype=np.uint8, copy=True)
    xfers = [o.Xfer(*x[:6]) for x in r["xfers"]]
    o.apply_transfers(store, xfers)                   # destinations start poisoned: a missed byte cannot pass
    received = {(x.dst, x.layer) for x in xfers if x.size > 0}
    tables = {(t["node"], t["layer"]): t for t in r["tables"]}
    checked = 0
    for node, layers in cfg.assignment.items():
        for layer in layers:
            assert (node, layer) in store, f"the oracle's store lacks ({node}, {layer}): the plan does not satisfy the assignment"
            want = store[(node, layer)]
            t = tables[(node, layer)]
            assert t["size"] == want.size, (node, layer)
            edge = min(4096, want.size)
            assert bytes.fromhex(t["head"]) == want[:edge].tobytes(), f"first bytes of layer {layer} at node {node} differ from the oracle"
            assert bytes.fromhex(t["tail"]) == want[want.size - edge:].tobytes(), f"last bytes of layer {layer} at node {node} differ"
            if (node, layer) not in received:
                continue                              # held from the start: nothing was accumulated into its table
            fused = np.frombuffer(bytes.fromhex(t["fused"]), dtype="<u8").reshape(-1, 2)
            assert np.array_equal(fused, o.checksum(want)), f"fused table of layer {layer} at node {node} differs from the oracle"
            assert t["received"] == want.size, (node, layer, t["received"])
            checked += 1
    return checked


def run_checked(text: str, mode: int, seed: int = 0, overrides=None, **opts) -> dict:
    """host.run_distribution + the oracle check above.  The report's `verified` must agree, but is not the criterion."""
    r = host.run_distribution(text, mode, seed=seed, dump_tables=True, **opts)
    r["oracle_checked"] = check_against_oracle(text, r, seed=seed, fill=opts.get("fill", o.FILL_PATTERN), overrides=overrides)
    assert r["verified"], "the library's own check disagrees with the oracle check that just passed"
    return r
