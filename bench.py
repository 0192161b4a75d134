#!/usr/bin/env python
"""bench.py — dissemination completion time and aggregate GB/s (BASELINE.json's metric).

Workload (config.workload): BASELINE config 2 — 8 nodes, 32 layers x 64 MiB, mode 1 (randomised peer
retransmission, seeded), leader holds all layers, node n is assigned stage 4n..4n+3 and initially holds its
ring-next peer's stage.  28 whole-layer transfers, D = 1 879 048 192 delivered bytes per step.  The 8 nodes are
mapped round-robin onto the N GPUs (node i -> rank i % N): at N=1 every transfer is an HBM-local copy, at N=8
every transfer crosses NVLink.  Total work is fixed as N grows ("strong").

A step = one pass of the hot path: launch the compiled schedule (one zeroing kernel + one copy+checksum kernel
per source GPU) and wait for completion.  `value` = D / (time of K steps / K), device-resident inputs.
`e2e` = the same through the C ABI with HOST buffers: every step uploads the source layers the plan reads from
pinned host memory, plans, compiles, launches, waits, and reads every destination's checksum table back.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--variant V]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_NODES, N_LAYERS, LAYER_BYTES = 8, 32, 67_108_864
MODE, SEED = 1, 0x5EED
METRIC = "dissemination_aggregate_throughput"
UNIT = "GB/s"
WORKLOAD = "config2: 8 nodes x 32 layers x 64 MiB, mode 1 (seeded random owner), ring-neighbour partial placement"


def config_text(n_nodes=N_NODES, n_layers=N_LAYERS, size=LAYER_BYTES) -> str:
    """The workload as reference-schema config JSON (cmd/config.go:14-45, nested form)."""
    per = n_layers // n_nodes
    stage = lambda n: range(per * n, per * n + per)
    bw = 900_000_000_000

    def node(i, leader, layers):
        nd = {"Id": i, "Addr": f"gpu:{i}", "NetworkBW": bw, "IsLeader": leader, "Sources": {"2": bw}}
        nd["InitialLayers"] = {"2": {str(l): {"LayerSize": size} for l in layers}} if layers else {}
        return nd

    nodes = [node(0, True, range(n_layers))]
    for n in range(1, n_nodes):
        nxt = n + 1 if n + 1 < n_nodes else 1
        nodes.append(node(n, False, stage(nxt) if n_nodes > 2 else []))
    return json.dumps({"Nodes": nodes, "Assignment": {str(n): {str(l): {} for l in stage(n)} for n in range(n_nodes)}})


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """SM clock + throttle reasons sampled DURING the timed region (NVML, ~1 kHz; nvidia-smi as a fallback)."""

    def __init__(self, cuda_index=0):
        self.cuda_index, self.rows, self._stop, self._t = cuda_index, [], threading.Event(), None
        self.nv = self.h = None
        try:
            import pynvml
            pynvml.nvmlInit()
            try:
                import torch
                uuid = str(torch.cuda.get_device_properties(cuda_index).uuid)
                if not uuid.startswith("GPU-"):
                    uuid = "GPU-" + uuid
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid.encode() if hasattr(uuid, "encode") else uuid)
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByIndex(cuda_index)
            self.nv = pynvml
        except Exception:
            self.nv = None

    def _sample_nvml(self):
        nv, h = self.nv, self.h
        sm = nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        try:
            r = nv.nvmlDeviceGetCurrentClocksEventReasons(h)
        except Exception:
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
        names = []
        for name, bit in (("hw_slowdown", 0x8), ("hw_thermal_slowdown", 0x40), ("sw_thermal_slowdown", 0x20),
                          ("sw_power_cap", 0x4)):
            if r & bit:
                names.append(name)
        self.rows.append((time.perf_counter(), sm, mx, names))

    def _sample_smi(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.cuda_index)],
                             capture_output=True, text=True, timeout=5).stdout.strip()
        if out:
            c = [v.strip() for v in out.split(",")]
            names = [n for n, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), c[2:6])
                     if v.lower().startswith("active")]
            self.rows.append((time.perf_counter(), int(float(c[0])), int(float(c[1])), names))

    def _run(self):
        while not self._stop.is_set():
            try:
                if self.nv is not None:
                    self._sample_nvml()
                else:
                    self._sample_smi()
            except Exception:
                pass
            self._stop.wait(0.001)

    def __enter__(self):
        self._t = threading.Thread(target=self._run, daemon=True)
        self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        self._t.join(timeout=6)

    def wait_first(self, timeout=3.0):
        end = time.perf_counter() + timeout
        while not self.rows and time.perf_counter() < end:
            time.sleep(0.005)

    def summary(self, t0=None, t1=None):
        """Samples taken between t0 and t1 (perf_counter), i.e. inside the timed region."""
        rows = [r for r in self.rows if (t0 is None or r[0] >= t0) and (t1 is None or r[0] <= t1)]
        sm = sorted(r[1] for r in rows)
        reasons = sorted({n for r in rows for n in r[3]})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max((r[2] for r in rows), default=None),
                "reasons": reasons, "samples": len(sm), "source": "nvml" if self.nv is not None else "nvidia-smi"}


# ------------------------------------------------------------------------------------------------------
# reference arm: the reference's own CPU path (C restatement of its TCP transport; Go is not installed)
# ------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world):
    if rank != 0:
        return
    from oracle import c_oracle as co
    from oracle import dissem_oracle as o
    cfg = o.parse_config(config_text())
    xfers = o.plan_mode1(cfg, seed=SEED, policy="random")
    layers = {}
    for x in xfers:
        if x.layer not in layers:
            layers[x.layer] = co.fill_bytes(o.FILL_PATTERN, x.layer, x.total, SEED)
    # calibrate, then bound each step so that the whole run ends within ~2 minutes
    t_cal, b_cal = co.tcp_run(N_NODES, 0, xfers[:4], [layers[x.layer] for x in xfers[:4]], base_port=21080)
    rate = b_cal / t_cal
    total_steps = args.steps + args.warmup
    budget_bytes = rate * 120.0 / max(total_steps, 1)
    n_take = max(1, min(len(xfers), int(budget_bytes // LAYER_BYTES)))
    sample = xfers[:n_take]
    srcs = [layers[x.layer] for x in sample]
    port = 21200
    for _ in range(args.warmup):
        co.tcp_run(N_NODES, 0, sample, srcs, base_port=port, verify=False)
        port += 16
    secs, moved = 0.0, 0
    for i in range(args.steps):
        s, b = co.tcp_run(N_NODES, 0, sample, srcs, base_port=port, verify=(i == 0))
        port += 16
        secs += s
        moved += b
    gbps = moved / secs / 1e9
    cores = os.cpu_count() or 1
    sample_desc = f"{n_take} of 28 transfers (x 64 MiB) per step over loopback TCP, one thread per connection"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gbps, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": {"workload": WORKLOAD, "sample": sample_desc},
        "cpu_baseline": {"value": gbps, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample_desc},
        "e2e": {"value": gbps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "C restatement of the reference's TCP layer path (oracle/oracle.c); the Go reference cannot be "
                "built here (no Go toolchain)"}))


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch

    import importlib

    import dissem_b200 as d
    multi = importlib.import_module("distributed-llm-dissemination_b200.multi")

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the engine has no CPU path")
    torch.cuda.set_device(local_rank)
    dist = None
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    text = config_text()
    plan = d.plan(text, MODE, seed=SEED, policy=0)
    xfers = d.parse_plan(plan)
    D = sum(x.size for x in xfers)
    owner = multi.owner_round_robin(world)
    eng = d.Engine(multi.device_of_node(N_NODES, owner, rank, local_rank))
    eng.set_option("variant", args.variant)
    eng.set_option("pull", 1 if args.pull == 1 else 0)
    pull = args.pull == 1
    if args.pull == 2:
        xfers = d.plan_directions(xfers, [owner(n) for n in range(N_NODES)])
    n_pulled = sum(1 for x in xfers if pull or (x.flags & 1))
    if args.tma_cfg is not None:
        eng.set_option("tma_cfg", args.tma_cfg)
    if args.reg_cfg is not None:
        eng.set_option("reg_cfg", args.reg_cfg)
    if args.ctas_per_sm is not None:
        eng.set_option("ctas_per_sm", args.ctas_per_sm)

    # the initial store (CreateLayers, outside the clock as in the reference) + poisoned destinations
    for nd in plan["nodes"]:
        if owner(nd["id"]) != rank:
            continue
        for l, size in nd["layers"].items():
            eng.alloc(nd["id"], int(l), size)
            eng.fill(nd["id"], int(l), d.FILL_PATTERN, SEED)
    for x in xfers:
        if owner(x.dst_node) == rank and not eng.has(x.dst_node, x.layer):
            eng.alloc(x.dst_node, x.layer, x.total)
            eng.fill(x.dst_node, x.layer, d.FILL_POISON)
    if dist is not None:
        # one process per GPU: ship the destination buffers' handles (the announce would carry them)
        mine = {(n, l): eng.export(n, l) for (n, l, _t) in multi.exports_needed(xfers, owner, rank, pull)}
        handles = multi.exchange_handles(dist, world, rank, mine)
        for (n, l, total) in multi.imports_needed(xfers, owner, rank, pull):
            eng.import_(n, l, total, handles[(n, l)])
    my_xfers = multi.my_transfers(xfers, owner, rank, pull)
    my_bytes = sum(x.size for x in my_xfers)
    sched = eng.compile(xfers)
    units, launches_per_step, sched_bytes = eng.sched_info(sched)
    assert sched_bytes == my_bytes

    # ---- device-resident throughput -------------------------------------------------------------------
    with ClockSampler(local_rank) as clocks:          # started before the warm-up so that NVML is warm; filtered below
        for _ in range(max(args.warmup, 3)):
            eng.wait(eng.launch(sched))
        clocks.wait_first()
        barrier()
        t0 = time.perf_counter()
        tickets = [eng.launch(sched) for _ in range(args.steps)]
        times = [eng.wait_detail(t) for t in tickets]
        barrier()
        t1 = time.perf_counter()
        wall = t1 - t0
        t_region0 = t0                      # (t0 is reused by the later legs)
        clock_window = "timed region"
        if clocks.summary(t0, t1)["samples"] < 3:
            # the timed region was too short for NVML (a few ms per query): keep the identical load running, outside the
            # timing, until a few samples exist, and say so
            t_more = time.perf_counter() + 0.25
            while time.perf_counter() < t_more:
                eng.wait(eng.launch(sched))
            clock_window = "timed region + the same launches continued for 0.25 s (region shorter than NVML's sampling period)"
        t_clk = time.perf_counter()
    dev_ms_total = sum(t[0] for t in times)
    copy_ms_avg = sum(t[1] for t in times) / max(len(times), 1)
    t_all = torch.tensor([wall, dev_ms_total], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t_all, op=dist.ReduceOp.MAX)
    wall_max, dev_ms_max = t_all.tolist()
    # K steps between barrier + synchronize on both sides, max over ranks (the launches are enqueued back to back on
    # one stream per GPU; the sum of their CUDA-event times is reported beside it as device_ms_per_step)
    step_s = wall_max / args.steps
    dev_step_s = dev_ms_max / 1e3 / args.steps

    # ---- correctness of what was just timed (outside the clock) ------------------------------------------
    ok = True
    for x in xfers:
        if owner(x.dst_node) != rank:
            continue
        fused = eng.checksums(x.dst_node, x.layer)
        ok = ok and np.array_equal(fused, eng.verify(x.dst_node, x.layer))
    if dist is not None:
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        ok = bool(flag.item())
    if not ok:
        raise SystemExit("bench.py: delivered layers failed verification")

    # ---- end to end through the C ABI with host buffers ---------------------------------------------------
    import ctypes
    L = d.lib()
    src_pairs = sorted({(x.src_node, x.layer) for x in xfers if owner(x.src_node) == rank})
    pinned = {}
    for (node, layer) in src_pairs:
        p = ctypes.c_void_p()
        assert L.dissem_host_alloc(ctypes.byref(p), LAYER_BYTES) == 0
        pinned[(node, layer)] = p
        data = eng.read(node, layer)                      # the host copy of the layer the node "loaded"
        ctypes.memmove(p.value, data.ctypes.data, LAYER_BYTES)
    my_dsts = [(x.dst_node, x.layer) for x in xfers if owner(x.dst_node) == rank]
    h2d = len(src_pairs) * LAYER_BYTES
    d2h = sum(((LAYER_BYTES + d.CHUNK_BYTES - 1) // d.CHUNK_BYTES) * 16 for _ in my_dsts)
    e2e_steps = max(2, min(args.steps, 5))

    def e2e_step():
        for (node, layer), p in pinned.items():
            eng.write(node, layer, None, ptr=p.value, size=LAYER_BYTES)
        if n_pulled and dist is not None:      # remote readers must not start before the upload has landed
            barrier()
        pl = d.parse_plan(d.plan(text, MODE, seed=SEED, policy=0))
        if args.pull == 2:
            pl = d.plan_directions(pl, [owner(n) for n in range(N_NODES)])
        t = eng.submit(pl)
        eng.wait(t)
        return [eng.checksums(n, l) for (n, l) in my_dsts]

    if args.no_e2e:
        e2e_steps = 0
    else:
        e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        e2e_step()
    barrier()
    e2e_wall = time.perf_counter() - t0
    te = torch.tensor([e2e_wall], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_step_s = te.item() / max(e2e_steps, 1)
    tot = torch.tensor([float(h2d), float(d2h)], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(tot, op=dist.ReduceOp.SUM)
    h2d_all, d2h_all = [int(v) for v in tot.tolist()]
    # ---- the same end to end, with the source layers LEFT in pinned host memory (dissem_layer_attach_host): the copy
    #      kernel reads them over PCIe, upload and dissemination are one pass.  One GPU only (reported beside e2e). ----
    e2e_direct = None
    if world == 1 and not args.no_e2e:
        eng2 = d.Engine([0] * N_NODES)
        for (node, layer), p in pinned.items():
            eng2.attach_host(node, layer, p.value, LAYER_BYTES)
        for x in xfers:
            if not eng2.has(x.dst_node, x.layer):
                eng2.alloc(x.dst_node, x.layer, x.total)

        def direct_step():
            pl = d.parse_plan(d.plan(text, MODE, seed=SEED, policy=0))
            eng2.wait(eng2.submit(pl))
            return [eng2.checksums(n, l) for (n, l) in my_dsts]

        got = direct_step()
        same = all(np.array_equal(a, eng.checksums(n, l)) for a, (n, l) in zip(got, my_dsts))
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            direct_step()
        torch.cuda.synchronize()
        dt = (time.perf_counter() - t0) / e2e_steps
        e2e_direct = {"value": D / dt / 1e9, "unit": UNIT, "ms_per_step": dt * 1e3, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "tables_match_staged_run": bool(same),
                      "what": "source layers attached in pinned host memory -> dissem_plan_json -> dissem_schedule_submit (kernel reads "
                              "host memory over PCIe) -> dissem_wait -> dissem_layer_checksums"}
        eng2.close()
    for p in pinned.values():
        L.dissem_host_free(p)

    # ---- CPU baseline on rank 0 (bounded sample of the same workload) -------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
      try:
          from oracle import c_oracle as co
          from oracle import dissem_oracle as o
          oxs = o.plan_mode1(o.parse_config(text), seed=SEED, policy="random")
          lay = {}
          for x in oxs:
              if x.layer not in lay:
                  lay[x.layer] = co.fill_bytes(o.FILL_PATTERN, x.layer, x.total, SEED)
          secs = moved = 0
          reps = 0
          t_begin = time.perf_counter()
          while reps < 400 and time.perf_counter() - t_begin < 12.0:      # about 10 s of CPU work
              s, b = co.tcp_run(N_NODES, 0, oxs, [lay[x.layer] for x in oxs], base_port=22080 + 16 * (reps % 64), verify=(reps == 0))
              secs += s
              moved += b
              reps += 1
          cpu = {"value": moved / secs / 1e9, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                 "sample": f"{reps} x the full workload (28 transfers x 64 MiB) over loopback TCP, one thread per connection; "
                           f"completion {secs / reps * 1e3:.1f} ms per pass"}
      except Exception as e:                       # the baseline is reported, never allowed to sink the bench line
          cpu = {"value": None, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port", "sample": f"failed: {e}"}

    if rank == 0:
        peak, peak_src = peaks()
        local = world == 1
        # algorithmic bytes of the dominant kernel: local copy = 2*S HBM bytes per delivered byte
        alg_bytes = 2 * my_bytes if local else my_bytes
        achieved = alg_bytes / (copy_ms_avg / 1e3) / 1e9 if copy_ms_avg > 0 else 0.0
        kernel_name = "copy_csum_tma_kernel" if args.variant == 1 else "copy_csum_kernel"
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(kernel_name + "_dram_bytes_per_launch")
            except Exception:
                traffic = None
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": kernel_name, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": copy_ms_avg}
        if not local:
            # NVLink roofline of the placement: T* = max over GPUs of max(egress, ingress) / link bandwidth, plus that
            # GPU's HBM-local copies; the measured 770 GB/s peer-copy figure is the denominator
            eg, ing, loc = multi.link_bytes(xfers, owner, world)
            link = 770.0
            t_star = max(max(e, i) / (link * 1e9) + 2 * l / (peak * 1e9) for e, i, l in zip(eg, ing, loc))
            busiest = max(max(e, i) for e, i in zip(eg, ing))
            roof = {"bound": "nvlink", "achieved": busiest / dev_step_s / 1e9, "peak": link, "unit": "GB/s",
                    "frac": t_star / dev_step_s, "traffic": None, "kernel": kernel_name,
                    "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                    "algorithmic_bytes_per_launch": busiest, "launch_ms": dev_step_s * 1e3,
                    "egress_bytes": eg, "ingress_bytes": ing, "local_bytes": loc, "t_star_ms": t_star * 1e3,
                    "frac_of_nominal_900": (t_star * link / 900.0) / dev_step_s,
                    "note": "achieved = bytes through the busiest GPU port per step / device time of a step (max over ranks)"}
        line = {
            "metric": METRIC, "value": D / step_s / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u8", "data": "synthetic",
            "config": {"workload": WORKLOAD, "delivered_bytes_per_step": D, "nodes_per_gpu": N_NODES // world if world <= 8 else 1,
                       "l2": "inputs (3.76 GB touched per step) far exceed the 126 MB L2; no flush needed",
                       "kernel_variant": "tma" if args.variant == 1 else "register", "mode": {0: "push", 1: "pull", 2: "auto"}[args.pull], "transfers_pulled": n_pulled, "units_per_launch": units,
                       "scaling_note": "the same 8-node plan on N GPUs: at N=1 every transfer is an HBM-local copy; from N=2 on the "
                                       "transfers cross NVLink and the busiest port bounds the step, so the value is not expected "
                                       "to grow with N (see roofline.frac for how close each N is to its own bound)"},
            "device_ms_per_step": dev_step_s * 1e3,
            "completion_time_ms": step_s * 1e3,
            "e2e": None if args.no_e2e else {"value": D / e2e_step_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
                    "ms_per_step": e2e_step_s * 1e3, "steps": e2e_steps,
                    "what": "pinned host layers -> dissem_layer_write -> dissem_plan_json -> dissem_schedule_submit -> "
                            "dissem_wait -> dissem_layer_checksums"},
            "e2e_host_sources": e2e_direct,
            "gpu_launches": launches_per_step * args.steps,
            "clocks": dict(clocks.summary(t_region0, t_clk), window=clock_window),
            "roofline": roof,
            "cpu_baseline": cpu,
            "verified": ok,
        }
        print(json.dumps(line))
    eng.sched_free(sched)
    eng.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", type=int, default=1, help="0 register path, 1 TMA bulk-staged path")
    ap.add_argument("--reg-cfg", type=int, default=None)
    ap.add_argument("--tma-cfg", type=int, default=None)
    ap.add_argument("--ctas-per-sm", type=int, default=None)
    ap.add_argument("--pull", type=int, default=2, help="0 push everything, 1 pull everything, 2 per-transfer choice "
                    "by dissem_plan_directions (default)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the end-to-end leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world == 1 and args.gpus > 1:
        # launched without torchrun: re-exec under it so that there is one process per GPU
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29517"), os.path.abspath(__file__)]
        cmd += sys.argv[1:]
        raise SystemExit(subprocess.call(cmd))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
