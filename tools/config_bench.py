"""Completion time of the BASELINE.json configurations under modes 0 / 1 / 3, ONE process over all visible GPUs
(node i on GPU i % n), device-resident layers, plans from the C++ planners (dissem_plan_json), direction choice on.
One JSON line per (config, mode).  Correctness is checked on the GPU after each case (fused table == re-read of
destination == re-read of source).

    python tools/config_bench.py [--configs readme,ring64,seed64,llama8b] [--reps 5]
"""
import argparse
import json
import sys

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import torch

import dissem_b200 as d

BW = 900_000_000_000


def node(i, leader, layers, size):
    nd = {"Id": i, "Addr": f"gpu:{i}", "NetworkBW": BW, "IsLeader": leader, "Sources": {"2": BW}}
    nd["InitialLayers"] = {"2": {str(l): {"LayerSize": size} for l in layers}} if layers else {}
    return nd


def stage_config(n_nodes, n_layers, size, ring):
    per = n_layers // n_nodes
    stage = lambda n: range(per * n, per * n + per)
    nodes = [node(0, True, range(n_layers), size)]
    for n in range(1, n_nodes):
        nxt = n + 1 if n + 1 < n_nodes else 1
        nodes.append(node(n, False, stage(nxt) if ring else [], size))
    return json.dumps({"Nodes": nodes, "Assignment": {str(n): {str(l): {} for l in stage(n)} for n in range(n_nodes)}})


def readme_config():
    return json.dumps({
        "Nodes": [{"Id": 0, "Addr": ":8080", "IsLeader": True, "NetworkBW": BW, "InitialLayers": {"1": {}, "3": {}}},
                  {"Id": 1, "Addr": ":8081", "IsLeader": False, "NetworkBW": BW, "InitialLayers": {"1": {}}},
                  {"Id": 2, "Addr": ":8082", "IsLeader": False, "NetworkBW": BW, "InitialLayers": {}},
                  {"Id": 3, "Addr": ":8083", "IsLeader": False, "NetworkBW": BW, "InitialLayers": {"3": {}}}],
        "Assignment": {"1": {"1": {}}, "2": {"1": {}, "3": {}}, "3": {"3": {}}}, "LayerSize": 1048576})


def reference_conf_json():
    """The reference's shipped conf/config.json (nested schema): nodes 0-6 each hold layers 0-7 of 10 930 691 768 B under
    source "1" (209 715 200 B/s), node 7 holds nothing and is assigned all eight; NetworkBW 1 562 500 000 everywhere."""
    size = 10_930_691_768
    nodes = []
    for i in range(8):
        nd = {"Id": i, "Addr": f":{8080 + i}", "NetworkBW": 1_562_500_000, "IsLeader": i == 0}
        if i < 7:
            nd["Sources"] = {"0": 16_257_500, "1": 209_715_200}
            nd["InitialLayers"] = {"1": {str(l): {"LayerSize": size} for l in range(8)}}
        else:
            nd["InitialLayers"] = {}
        nodes.append(nd)
    return json.dumps({"Nodes": nodes, "Assignment": {"7": {str(l): {} for l in range(8)}}})


CONFIGS = {
    "refconf": ("the reference's conf/config.json: 8 x 10 930 691 768 B from 7 holders to node 7 (its flow plan: t = 60 s)", reference_conf_json),
    "readme": ("config 1: README JSON, 4 nodes, 1 MiB layers", readme_config),
    "ring64": ("config 2: 8 x 32 x 64 MiB, ring-neighbour partial placement", lambda: stage_config(8, 32, 67_108_864, True)),
    "seed64": ("config 5: 8 x 32 x 64 MiB, one seed GPU", lambda: stage_config(8, 32, 67_108_864, False)),
    "llama8b": ("config 3: Llama-3-8B-shaped, 32 x 436 224 000 B, leader seeds all", lambda: stage_config(8, 32, 436_224_000, False)),
    "llama8b_218": ("config 3 (literal 218 MB): 32 x 218 112 000 B", lambda: stage_config(8, 32, 218_112_000, False)),
}


def run(name, desc, text, mode, policy, n_gpu, reps):
    flags = 1 if name == "readme" else 0
    try:
        plan = d.plan(text, mode, seed=0x5EED, policy=policy, time_scale=1 if name == "refconf" else 1_000_000, flags=flags)
    except d.DissemError as e:
        print(json.dumps({"config": name, "mode": mode, "error": str(e)}), flush=True)
        return
    n_nodes = max(nd["id"] for nd in plan["nodes"]) + 1
    devs = [i % n_gpu for i in range(n_nodes)]
    xfers = d.plan_directions(d.parse_plan(plan), devs)
    with d.Engine(devs) as eng:
        for nd in plan["nodes"]:
            for l, size in nd["layers"].items():
                eng.alloc(nd["id"], int(l), size)
                eng.fill(nd["id"], int(l), d.FILL_PATTERN, 0x5EED)
        for x in xfers:
            if not eng.has(x.dst_node, x.layer):
                eng.alloc(x.dst_node, x.layer, x.total)
                eng.fill(x.dst_node, x.layer, d.FILL_POISON)
        s = eng.compile(xfers)
        for _ in range(3):
            eng.wait(eng.launch(s))
        times = [eng.wait_detail(eng.launch(s))[0] for _ in range(reps)]
        ok = True
        seen = set()
        for x in xfers:
            if (x.dst_node, x.layer) in seen:
                continue
            seen.add((x.dst_node, x.layer))
            if len(seen) > 12:
                break
            ref = eng.verify(x.src_node, x.layer)
            ok = ok and bool((eng.checksums(x.dst_node, x.layer) == ref).all() and (eng.verify(x.dst_node, x.layer) == ref).all())
        D = sum(x.size for x in xfers)
        eg, ing = [0] * n_gpu, [0] * n_gpu
        for x in xfers:
            if devs[x.src_node] != devs[x.dst_node]:
                eg[devs[x.src_node]] += x.size
                ing[devs[x.dst_node]] += x.size
        best = min(times)
        print(json.dumps({"config": name, "what": desc, "mode": mode, "policy": ["random", "balanced"][policy], "n_gpus": n_gpu,
                          "transfers": len(xfers), "pulled": sum(1 for x in xfers if x.flags & 1), "delivered_bytes": D,
                          "plan_t": plan["t"], "completion_ms_best": best, "completion_ms_avg": sum(times) / len(times),
                          "aggregate_GBps": D / best / 1e6 if best > 0 else None, "busiest_port_bytes": max(eg + ing),
                          "busiest_port_GBps": max(eg + ing) / best / 1e6 if best > 0 else None, "verified": ok}), flush=True)
        eng.sched_free(s)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--configs", default="readme,ring64,seed64,llama8b")
    ap.add_argument("--reps", type=int, default=5)
    ap.add_argument("--gpus", type=int, default=0)
    args = ap.parse_args()
    n_gpu = args.gpus or min(torch.cuda.device_count(), 8)
    for name in args.configs.split(","):
        desc, make = CONFIGS[name]
        text = make()
        for mode, policy in ((0, 0), (1, 0), (1, 1), (3, 0)):
            run(name, desc, text, mode, policy, n_gpu, args.reps)


if __name__ == "__main__":
    main()
