"""Where a small Transport.Send spends its time (config 1 of BASELINE.json: README JSON, 1 MiB layers).
Prints one JSON line per measurement.   python tools/latency_probe.py [--size 1048576] [--reps 200]"""
import argparse
import importlib
import json
import sys
import time

sys.path.insert(0, __file__.rsplit("/tools/", 1)[0])
import dissem_b200 as d

host = importlib.import_module("distributed-llm-dissemination_b200.host")


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", type=int, default=1 << 20)
    ap.add_argument("--reps", type=int, default=200)
    args = ap.parse_args()
    size, reps = args.size, args.reps
    with d.Engine([0, 0]) as eng:
        eng.alloc(0, 0, size)
        eng.fill(0, 0, d.FILL_PATTERN, 1)
        eng.alloc(1, 0, size)
        eng.wait(eng.send(0, 1, 0, 0, size, size))
        t0 = time.perf_counter()
        for _ in range(reps):
            eng.wait(eng.send(0, 1, 0, 0, size, size))
        send_wait = (time.perf_counter() - t0) / reps
        t0 = time.perf_counter()
        for i in range(reps):
            eng.alloc(1, 1000 + i, size)
        alloc = (time.perf_counter() - t0) / reps
        t0 = time.perf_counter()
        for i in range(reps):
            eng.free(1, 1000 + i)
        free = (time.perf_counter() - t0) / reps
        eng.reserve(1, size, reps)
        t0 = time.perf_counter()
        for i in range(reps):
            eng.alloc(1, 5000 + i, size)
        alloc_reserved = (time.perf_counter() - t0) / reps
        print(json.dumps({"what": "engine calls, one layer", "layer_bytes": size, "send+wait_us": send_wait * 1e6,
                          "layer_alloc_us": alloc * 1e6, "layer_alloc_from_reserve_us": alloc_reserved * 1e6,
                          "layer_free_us": free * 1e6}), flush=True)
    text = json.dumps({          # the README's JSON (readme.md:17-64), flat schema
        "Nodes": [{"Id": 0, "Addr": ":8080", "IsLeader": True, "InitialLayers": {"1": {}, "3": {}}},
                  {"Id": 1, "Addr": ":8081", "IsLeader": False, "InitialLayers": {"1": {}}},
                  {"Id": 2, "Addr": ":8082", "IsLeader": False, "InitialLayers": {}},
                  {"Id": 3, "Addr": ":8083", "IsLeader": False, "InitialLayers": {"3": {}}}],
        "Assignment": {"1": {"1": {}}, "2": {"1": {}, "3": {}}, "3": {"3": {}}}, "LayerSize": size})
    for mode in (0, 1, 2):                      # mode 3 cannot plan the README JSON: it names no NetworkBW
        for batched in (False, True):
            for reserve in (0, 1):
                best = 1e9
                for _ in range(3):
                    r = host.run_distribution(text, mode, batched=batched, seed=1, time_scale=1, flags=3, verify=False, reserve=reserve)
                    best = min(best, r["time_to_deliver_ms"])
                print(json.dumps({"what": "README config through the roles", "mode": mode, "batched": batched, "reserve": reserve,
                                  "time_to_deliver_ms_best_of_3": best, "layer_msgs": r["layer_msgs"], "control_msgs": r["control_msgs"]}), flush=True)


if __name__ == "__main__":
    main()
