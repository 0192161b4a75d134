"""bench.py prints ONE JSON line with the keys the driver reads.  The reference arm runs on the CPU (here too); our arm
needs a GPU."""
import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BASE_KEYS = {"metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config", "e2e", "gpu_launches", "cpu_baseline"}


def _run(*args, timeout=600):
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), *args], capture_output=True, text=True,
                         timeout=timeout, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, out.stdout[-2000:]
    return json.loads(lines[0])


def test_reference_arm_line():
    j = _run("--impl", "reference", "--steps", "1", "--warmup", "0")
    assert BASE_KEYS <= set(j) and j["impl"] == "reference"
    assert j["metric"] == "dissemination_aggregate_throughput" and j["unit"] == "GB/s" and j["higher_is_better"] is True
    assert j["value"] > 0 and j["dtype"] == "u8" and j["vs_baseline"] is None and "workload" in j["config"]
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1 and j["cpu_baseline"]["value"] == j["value"]
    assert j["e2e"] == {"value": j["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["gpu_launches"] == 0


@pytest.mark.gpu
def test_our_arm_line():
    j = _run("--steps", "5", "--warmup", "3")
    assert BASE_KEYS | {"roofline", "clocks", "verified"} <= set(j) and "impl" not in j
    assert j["n_gpus"] == 1 and j["steps"] == 5 and j["warmup"] >= 3 and j["scaling"] == "strong" and j["verified"] is True
    assert j["value"] > 100 and abs(j["ms_per_step"] - j["config"]["delivered_bytes_per_step"] / j["value"] / 1e6) < 1e-6
    r = j["roofline"]
    assert r["bound"] == "hbm" and r["unit"] == "GB/s" and abs(r["frac"] - r["achieved"] / r["peak"]) < 1e-9
    assert r["algorithmic_bytes_per_launch"] == 2 * j["config"]["delivered_bytes_per_step"]
    e = j["e2e"]
    assert e["h2d_bytes_per_step"] == j["config"]["delivered_bytes_per_step"] and e["d2h_bytes_per_step"] > 0
    assert 0 < e["value"] < j["value"]                      # the end-to-end number is its own measurement
    assert j["gpu_launches"] == 1 * 5                     # whole layers, units == checksum chunks: sums are stored, no zeroing kernel
    assert j["completion_time_ms"] > 0 and j["completion_device_ms"] > 0 and j["config"]["workload"].startswith("config3")
    assert j["cpu_baseline"]["kind"] == "port" and j["cpu_baseline"]["cores"] >= 1
    c = j["clocks"]
    assert not ({"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"} & set(c["reasons"]))
    ref = _run("--impl", "reference", "--steps", "1", "--warmup", "0")
    assert ref["config"] == j["config"]                     # the two arms describe the same workload, key for key


def test_go_probe_reports_why():
    """the reference arm probes for a Go toolchain + offline module cache at run time and says what it
    found.  On this image (and the GPU boxes) there is no `go`: the probe must say so, never raise, never download."""
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    p = bench.probe_go()
    assert isinstance(p, dict) and "usable" in p
    if not p["usable"]:
        assert p["why"]
    else:
        assert os.path.exists(p["binary"])
    j = _run("--impl", "reference", "--steps", "1", "--warmup", "0")
    assert "Go probe:" in j["note"]


def test_best_sender_t_star_matches_the_hand_computed_port_loads():
    """bench.py's roofline denominator at N >= 2: the busiest NVLink port under the BEST sender choice for config 2 with node i
    on GPU i % N: 268 MB at 8 GPUs, 537 MB at 4, 805 MB at 2 (nodes served inside their own GPU cost no port),
    divided by the measured 770 GB/s."""
    sys.path.insert(0, ROOT)
    import importlib
    bench = importlib.import_module("bench")
    import dissem_b200 as d
    text = bench.config_text(bench.CFG2["size"], bench.CFG2["ring"])
    plan = d.plan(text, bench.PLAN_MODE, time_scale=bench.TIME_SCALE)
    stage = 4 * bench.CFG2["size"]
    for world, want in ((8, stage), (4, 2 * stage), (2, 3 * stage)):
        gpus = [n % world for n in range(bench.N_NODES)]
        t_star, port = bench.best_sender_t_star(plan["nodes"], plan["assignment"], gpus, bench.CFG2["size"], world)
        assert port == want, (world, port)
        assert abs(t_star - want / 770e9) < 1e-12
    # one GPU: everything is local, nothing crosses a port
    assert bench.best_sender_t_star(plan["nodes"], plan["assignment"], [0] * 8, bench.CFG2["size"], 1) == (0.0, 0)
