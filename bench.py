#!/usr/bin/env python
"""bench.py — dissemination completion time and aggregate GB/s (BASELINE.json's metric).

Workload (config.workload):
  * 1 GPU   BASELINE config 3, the largest configuration that fits one GPU: Llama-3-8B-shaped, 32 layers x 436 224 000 B,
            node n is assigned stage 4n..4n+3, the leader seeds all.  28 whole-layer transfers, D = 12 214 272 000 bytes
            per step, every one an HBM-local copy (26 GB resident).
  * N >= 2  BASELINE config 2: 8 nodes, 32 layers x 64 MiB, ring-neighbour partial placement, node i on GPU i % N,
            D = 1 879 048 192 bytes per step, executed with the engine's best schedule for the placement: the flow.go plan
            (mode 3, 1 us ticks) -> dissem_plan_localize -> dissem_plan_optimize.  The seeded mode-1 pick of round 1 is
            timed beside it as an extra.  The roofline is T* under the BEST sender choice, not the executed
            plan's own busiest port.
  * N == 8  additionally a `north_star` block: the full Llama-3-70B-shaped assignment (80 x 1 711 308 800 B, fan-out 2-4),
            balanced and one-seed placements — ours, NVLS multicast, NCCL send/recv and NCCL broadcast renditions.

A step = one pass of the hot path: launch the compiled schedule (one zeroing kernel + one copy+checksum kernel per
executing GPU and round) and wait for completion.  `value` = D / (time of K pipelined steps / K), device-resident
inputs; `completion_time_ms` = median of single-shot steps (barrier, launch, wait, barrier).  `e2e` = the same through
the C ABI with HOST buffers: every step uploads the byte ranges the plan reads from pinned host memory, plans, compiles,
launches, waits, and reads every destination's checksum table back.  Everything delivered is checked against the CPU
oracle (tables of the seeded fill, spot bytes) after the timed regions.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N ...
"""
import argparse
import json
import os
import shutil
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

N_NODES, N_LAYERS = 8, 32
SEED = 0x5EED
PLAN_MODE, TIME_SCALE = 3, 1_000_000
METRIC = "dissemination_aggregate_throughput"
UNIT = "GB/s"
LINK_GBPS, LINK_NOMINAL_GBPS = 770.0, 900.0       # measured peer copy per direction (B200_PROFILING.md); nominal NVLink 5

CFG3 = {"key": "config3", "size": 436_224_000, "ring": False,
        "name": "config3: Llama-3-8B-shaped, 8 nodes x 32 layers x 436224000 B, node n <- layers 4n..4n+3, leader seeds all "
                "(28 transfers, flow.go plan)"}
CFG2 = {"key": "config2", "size": 67_108_864, "ring": True,
        "name": "config2: 8 nodes x 32 layers x 64 MiB, ring-neighbour partial placement, flow.go plan (mode 3, 1 us ticks) "
                "-> dissem_plan_localize -> dissem_plan_optimize"}


def workload_for(world: int) -> dict:
    return CFG3 if world == 1 else CFG2


def config_text(size: int, ring: bool, n_nodes=N_NODES, n_layers=N_LAYERS) -> str:
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
        nodes.append(node(n, False, stage(nxt) if ring else []))
    return json.dumps({"Nodes": nodes, "Assignment": {str(n): {str(l): {} for l in stage(n)} for n in range(n_nodes)}})


def bench_config(w: dict, world: int, D: int) -> dict:
    """The `config` object, identical in both arms."""
    return {"workload": w["name"], "delivered_bytes_per_step": D, "layer_bytes": w["size"], "nodes": N_NODES, "layers": N_LAYERS,
            "nodes_per_gpu": N_NODES // world if N_NODES % world == 0 else None,
            "l2": "inputs touched per step (%.2f GB) far exceed the 126 MB L2; no flush needed" % (2 * D / 1e9),
            "scaling_note": "1 GPU runs config 3 (HBM-local copies); N >= 2 run config 2 across NVLink with total work fixed, "
                            "so the value is bounded by the placement's busiest port, not by N (see roofline)"}


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def mem_available_gb() -> float:
    try:
        for line in open("/proc/meminfo"):
            if line.startswith("MemAvailable:"):
                return int(line.split()[1]) / 1e6
    except Exception:
        pass
    return 0.0


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
# reference arm: the reference's own CPU path.  The Go distributor itself when a Go toolchain and its modules
# are present (probed below); otherwise the C restatement of its TCP transport (oracle/oracle.c).
# ------------------------------------------------------------------------------------------------------
def probe_go() -> dict:
    """can the real Go distributor be built on this box?  Needs `go` and the three modules of go.mod in
    an offline module cache, and the reference tree (absent on the GPU box).  Never installs or downloads anything."""
    go = shutil.which("go")
    ref = "/root/reference"
    if not go:
        return {"go": None, "usable": False, "why": "no `go` on PATH"}
    if not os.path.exists(os.path.join(ref, "go.mod")):
        return {"go": go, "usable": False, "why": "reference tree not on this box"}
    try:
        env = dict(os.environ, GOFLAGS="-mod=mod", GOPROXY="off", GOTOOLCHAIN="local")
        out_dir = os.path.join(ROOT, "oracle", "_ref")
        os.makedirs(out_dir, exist_ok=True)
        r = subprocess.run([go, "build", "-o", os.path.join(out_dir, "distributor"), "./cmd"], cwd=ref, env=env,
                           capture_output=True, text=True, timeout=120)
        if r.returncode != 0:
            return {"go": go, "usable": False, "why": "go build failed offline: " + (r.stderr or r.stdout)[-160:].replace("\n", " ")}
        return {"go": go, "usable": True, "binary": os.path.join(out_dir, "distributor")}
    except Exception as e:                                     # noqa: BLE001
        return {"go": go, "usable": False, "why": repr(e)[:160]}


class CpuTcp:
    """The oracle's loopback-TCP restatement on the host cores over the workload's transfer list (the oracle's own
    flow.go plan).  The sample is bounded by host memory (sources + freshly allocated receive buffers ~ 2 x its bytes);
    a pass moves every transfer of the sample once."""

    def __init__(self, w: dict):
        from concurrent.futures import ThreadPoolExecutor

        from oracle import c_oracle as co
        from oracle import dissem_oracle as o
        self.co = co
        cfg = o.parse_config(config_text(w["size"], w["ring"]))
        _t, xfers, _self_jobs = o.plan_mode3(cfg, time_scale=TIME_SCALE)
        self.total = sum(x.size for x in xfers)
        budget = max(2e9, min(mem_available_gb() * 1e9 * 0.25, 16e9))
        self.sample, self.bytes = [], 0
        for x in xfers:
            if self.sample and self.bytes + x.size > budget:
                break
            self.sample.append(x)
            self.bytes += x.size
        ids = sorted({x.layer for x in self.sample})
        size = {x.layer: x.total for x in self.sample}
        with ThreadPoolExecutor(max_workers=min(len(ids), max(1, (os.cpu_count() or 2) // 2))) as ex:
            filled = list(ex.map(lambda l: co.fill_bytes(o.FILL_PATTERN, l, size[l], SEED), ids))
        layers = dict(zip(ids, filled))
        self.srcs = [layers[x.layer] for x in self.sample]
        self.n_all = len(xfers)
        self.passes = 0

    def run_pass(self, verify=False):
        """-> (seconds, bytes moved)"""
        s, b = self.co.tcp_run(N_NODES, 0, self.sample, self.srcs, base_port=21080 + 16 * (self.passes % 64), verify=verify)
        self.passes += 1
        return s, b

    def shrink(self, n_keep: int):
        """Keep the first n_keep transfers of the sample (a slow host: keeps the whole run within minutes)."""
        n_keep = max(1, min(n_keep, len(self.sample)))
        self.sample, self.srcs = self.sample[:n_keep], self.srcs[:n_keep]
        self.bytes = sum(x.size for x in self.sample)

    def describe(self):
        return (f"{len(self.sample)} of {self.n_all} transfers ({self.bytes / 1e9:.2f} of {self.total / 1e9:.2f} GB per pass) over "
                f"loopback TCP, one thread per connection, {self.passes} passes")


def run_reference(args, rank, world):
    if rank != 0:
        return
    w = workload_for(world)
    go = probe_go()
    tcp = CpuTcp(w)
    s0, _b0 = tcp.run_pass()                       # sizing pass (untimed): the whole run has to end within ~2 minutes
    total_passes = args.steps + args.warmup
    if s0 * total_passes > 120.0:
        tcp.shrink(int(len(tcp.sample) * 120.0 / (s0 * total_passes)))
    for _ in range(args.warmup):
        tcp.run_pass()
    secs = moved = 0.0
    for i in range(args.steps):                    # one step = one pass over the (memory-bounded) sample
        s_, b_ = tcp.run_pass(verify=(i == 0))
        secs += s_
        moved += b_
    gbps = moved / secs / 1e9
    cores = os.cpu_count() or 1
    desc = tcp.describe()
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": gbps, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic",
        "config": bench_config(w, world, tcp.total),
        "cpu_baseline": {"value": gbps, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": gbps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "C restatement of the reference's TCP layer path (oracle/oracle.c), one step = one pass over the sample. "
                "Go probe: " + (f"usable ({go['binary']}) but the multi-process harness is not wired; port timed"
                                if go.get("usable") else f"reference not runnable here ({go['why']})")}))


# ------------------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------------------
LINE = {}
_EMITTED = threading.Event()


def emit(rank):
    """Print the one JSON line (rank 0), once."""
    if _EMITTED.is_set():
        return
    _EMITTED.set()
    if rank == 0:
        print(json.dumps(LINE), flush=True)


def best_sender_t_star(plan_nodes, assignment, gpu_of_node, size, world):
    """T* under the best sender choice = the smallest busiest-port load C (bytes) for which every
    (destination, layer) pair that no node on the destination's GPU already holds can be supplied by the GPUs that hold
    the layer, with every GPU's egress and ingress <= C.  Small max-flow inside a bisection (pure Python; 28 demands)."""
    holders = {}
    for nd in plan_nodes:
        for l in nd["layers"]:
            holders.setdefault(int(l), set()).add(gpu_of_node[nd["id"]])
    demands = []                                   # (dst gpu, [holder gpus])
    for n, layers in assignment.items():
        g = gpu_of_node[int(n)]
        for l in layers:
            hs = holders.get(int(l), set())
            if g in hs:
                continue                           # served inside the GPU (or already in place)
            demands.append((g, sorted(hs)))
    if not demands:
        return 0.0, 0
    need = len(demands) * size

    def feasible(cap):
        # vertices: 0 src, 1..world egress, then demands, then ingress, sink
        n_d = len(demands)
        S, T = 0, 1 + world + n_d + world
        adj = [dict() for _ in range(T + 1)]

        def add(a, b, c):
            adj[a][b] = adj[a].get(b, 0) + c
            adj[b].setdefault(a, 0)

        for g in range(world):
            add(S, 1 + g, cap)
            add(1 + world + n_d + g, T, cap)
        for i, (g, hs) in enumerate(demands):
            for h in hs:
                add(1 + h, 1 + world + i, size)
            add(1 + world + i, 1 + world + n_d + g, size)
        flow = 0
        while True:
            parent = {S: None}
            queue = [S]
            for u in queue:
                for v, c in adj[u].items():
                    if c > 0 and v not in parent:
                        parent[v] = u
                        queue.append(v)
                if T in parent:
                    break
            if T not in parent:
                break
            push, v = float("inf"), T
            while parent[v] is not None:
                push = min(push, adj[parent[v]][v])
                v = parent[v]
            v = T
            while parent[v] is not None:
                adj[parent[v]][v] -= push
                adj[v][parent[v]] += push
                v = parent[v]
            flow += push
        return flow >= need

    lo, hi = 0, need
    while lo < hi:
        mid = (lo + hi) // 2
        if feasible(mid):
            hi = mid
        else:
            lo = mid + 1
    return lo / (LINK_GBPS * 1e9), lo


class Runner:
    """A plan on one engine.  One process: one schedule, rounds ordered on the device by events.  One process per GPU:
    a plan with rounds runs flag-driven when it can (multi.FlagRelaySchedule: all rounds in one schedule, hops wait on the
    device for the source copy's received counter, one reset + barrier before the dissemination), else round by round
    with a barrier between rounds."""

    def __init__(self, eng, multi, xfers, barrier, multi_process, owner=None, rank=0, allow_flags=True):
        self.eng, self.barrier, self.mp = eng, barrier, multi_process
        rounds = multi.split_rounds(xfers)
        self.flag = None
        if multi_process and len(rounds) > 1 and allow_flags and owner is not None and multi.flag_relay_ok(xfers):
            self.flag = multi.FlagRelaySchedule(eng, xfers, owner, rank)
            self.scheds = self.flag.scheds
        else:
            self.scheds = [eng.compile(r) for r in (rounds if multi_process else [list(xfers)])]
        info = [eng.sched_info(s) for s in self.scheds]
        self.units = sum(i[0] for i in info)
        self.launches = sum(i[1] for i in info)
        self.bytes = sum(i[2] for i in info)
        self.n_rounds = len(rounds)

    def describe(self):
        if self.flag is not None:
            return f"{self.n_rounds} rounds in one launch sequence, ordered on the device by received-counter waits"
        if len(self.scheds) > 1:
            return f"{len(self.scheds)} rounds, a barrier between rounds"
        return f"{self.n_rounds} round(s), one schedule"

    def step(self):
        """One dissemination, waited for.  -> (device ms, copy-kernel ms) of this rank."""
        if self.flag is not None:
            return self.flag.run(self.barrier)
        dev = copy = 0.0
        for s in self.scheds:
            ms, cms = self.eng.wait_detail(self.eng.launch(s))
            dev += ms
            copy += cms
            if len(self.scheds) > 1:
                self.barrier()
        return dev, copy

    def pipelined(self, k):
        """k disseminations enqueued back to back (single-round plans), then waited for."""
        if len(self.scheds) > 1 or self.flag is not None:
            return [self.step() for _ in range(k)]
        tickets = [self.eng.launch(self.scheds[0]) for _ in range(k)]
        return [self.eng.wait_detail(t) for t in tickets]

    def free(self):
        if self.flag is not None:
            self.flag.free()
        else:
            for s in self.scheds:
                self.eng.sched_free(s)
        self.scheds = []


def merged_ranges(ranges):
    out = []
    for a, b in sorted(ranges):
        if out and a <= out[-1][1]:
            out[-1][1] = max(out[-1][1], b)
        else:
            out.append([a, b])
    return [(a, b) for a, b in out]


def oracle_check(eng, co, o, dsts, size, expected, fill_seed=SEED, reads=True):
    """Fused tables (one read-back per GPU) and spot bytes of every destination in `dsts` against the CPU oracle."""
    import numpy as np
    if not dsts:
        return True
    got = eng.checksums_many(dsts)
    ok = all(np.array_equal(g, expected[l]) for g, (_n, l) in zip(got, dsts))
    if reads and ok:
        win = 65536
        for (n, l) in dsts:
            for off in sorted({0, max(0, (size // 2) // 8 * 8 - 8), max(0, (size - win) // 8 * 8)}):
                ln = min(win, size - off)
                if ln > 0 and not np.array_equal(eng.read(n, l, off, ln), co.fill_range(o.FILL_PATTERN, l, off, ln, fill_seed)):
                    return False
    return ok


def parity_leg(d, multi, co, o, torch, dist, rank, world, local_rank, barrier):
    """Small ragged layers through every path that crosses NVLink — push, pull, NVLS multicast, relay over rounds —
    destinations poisoned first, bytes and fused tables read back by the owning rank and compared with the C oracle."""
    import numpy as np
    size = 9 * d.CHUNK_BYTES + 4321
    seed = 0xFACE
    owner = lambda node: node
    res = {}

    def check(eng, pairs):
        ok = True
        for (n, l) in pairs:
            want = co.fill_bytes(o.FILL_PATTERN, l, size, seed)
            ok = ok and np.array_equal(eng.read(n, l), want) and np.array_equal(eng.checksums(n, l), co.checksum(want)) \
                and eng.received(n, l) == size
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        return bool(flag.item())

    nxt = lambda r: (r + 1) % world
    # ---- push and pull on IPC-mapped buffers: rank r's layer r goes to rank r+1 in three odd-cut fragments ----------
    cuts = [0, 70001, 5 * d.CHUNK_BYTES + 13, size]
    for name, flag in (("push", 0), ("pull", d.XFER_PULL)):
        eng = d.Engine(multi.device_of_node(world, owner, rank, local_rank))
        eng.alloc(rank, rank, size)
        eng.fill(rank, rank, d.FILL_PATTERN, seed)
        src_layer = (rank - 1) % world
        eng.alloc(rank, src_layer, size)
        eng.fill(rank, src_layer, d.FILL_POISON)
        xs = [d.Xfer(r, nxt(r), r, flag, cuts[i], cuts[i + 1] - cuts[i], size) for r in range(world) for i in range(3)]
        store = multi.SharedStore(eng, dist, world, rank, owner)
        store.share(xs)
        barrier()
        eng.wait(eng.submit(xs))
        barrier()
        res[name] = check(eng, [(rank, src_layer)])
        barrier()
        eng.close()
    # ---- relay over rounds: layer 100 from rank 0 along the chain 1, 2, ... in 3 pieces ------------------------------
    eng = d.Engine(multi.device_of_node(world, owner, rank, local_rank))
    eng.alloc(rank, 100, size)
    eng.fill(rank, 100, d.FILL_PATTERN if rank == 0 else d.FILL_POISON, seed)
    base = [d.Xfer(0, r, 100, 0, 0, size, size) for r in range(1, world)]
    xs = d.plan_directions(d.plan_relay(base, 3), list(range(world)))
    store = multi.SharedStore(eng, dist, world, rank, owner)
    store.share(xs)
    barrier()
    for name, flags_on in (("relay", True), ("relay_barrier_per_round", False)):
        if name != "relay":
            eng.fill(rank, 100, d.FILL_PATTERN if rank == 0 else d.FILL_POISON, seed)
            if rank != 0:
                eng.reset(rank, 100)
            barrier()
        run = Runner(eng, multi, xs, barrier, True, owner=owner, rank=rank, allow_flags=flags_on)
        run.step()
        barrier()
        timeouts = torch.tensor([eng.get_option("wait_timeouts")], device="cuda")
        dist.all_reduce(timeouts, op=dist.ReduceOp.SUM)
        res[name] = check(eng, [(rank, 100)] if rank != 0 else []) and int(timeouts.item()) == 0
        res[name + "_how"] = run.describe()
        run.free()
        barrier()
    eng.close()
    # ---- NVLS multicast: layer 200 from rank 0 to every other rank, shareable VMM layers, handles as descriptors -----
    size = 9 * d.CHUNK_BYTES + 4320                  # multimem.st moves 16-byte vectors: multicast needs a multiple of 16
    fx = multi.FdExchange(rank, "par" + os.environ.get("MASTER_PORT", "0"))
    barrier()
    eng = d.Engine(multi.device_of_node(world, owner, rank, local_rank))
    eng.set_option("alloc_vmm", 2)
    eng.alloc(rank, 200, size)
    eng.fill(rank, 200, d.FILL_PATTERN if rank == 0 else d.FILL_POISON, seed)
    xs = [d.Xfer(0, r, 200, 0, 0, size, size) for r in range(1, world)]
    store = multi.SharedStore(eng, dist, world, rank, owner, fx=fx)
    store.share(xs)
    mc_ok, why = multi.setup_mcast(eng, fx, dist, [(200, 0, list(range(1, world)))], owner, rank)
    if mc_ok:
        eng.set_option("mcast", 2 if world == 2 else 3)
        barrier()
        groups = 0
        if rank == 0:
            eng.wait(eng.submit(xs))
            groups = eng.get_option("mcast_groups")
        barrier()
        g = torch.tensor([groups], device="cuda")
        dist.all_reduce(g, op=dist.ReduceOp.MAX)
        res["mcast"] = check(eng, [(rank, 200)] if rank != 0 else []) and int(g.item()) >= 1
        res["mcast_groups"] = int(g.item())
    else:                                            # capability missing on this box: reported, not hidden
        res["mcast"] = None
        res["mcast_note"] = why[:200]
    barrier()
    eng.close()
    fx.close()
    return res


def north_star(d, multi, co, o, torch, dist, rank, world, local_rank, barrier, n_layers, size, reps, progress):
    """The full Llama-3-70B-shaped assignment on `world` GPUs (node r on rank r), balanced and one-seed placements:
    ours (dissem_plan_optimize), NVLS multicast, NCCL grouped send/recv and NCCL broadcast per layer."""
    import numpy as np
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import nccl_schedule
    owner = lambda node: node
    gpus = list(range(world))
    out = {"layers": n_layers, "layer_bytes": size, "n_gpus": world}
    # expected tables of the seeded fill, once: each rank computes its share on the host cores, then they are gathered
    mine = [l for l in range(n_layers) if l % world == rank]
    t0 = time.perf_counter()
    tabs = co.fill_checksum_many(o.FILL_PATTERN, mine, size, SEED, threads=max(1, (os.cpu_count() or 8) // world))
    everyone = [None] * world
    dist.all_gather_object(everyone, {int(k): v for k, v in tabs.items()})
    expected = {}
    for part in everyone:
        expected.update(part)
    out["oracle_tables_s"] = time.perf_counter() - t0
    nccl_groups = {}

    def timed(run):
        run.step()                                   # warm
        best, devs = 1e9, []
        for _ in range(reps):
            barrier()
            t0 = time.perf_counter()
            dev, _c = run.step()
            barrier()
            t = torch.tensor([time.perf_counter() - t0, dev / 1e3], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            best = min(best, t[0].item())
            devs.append(t[1].item())
        return best * 1e3, min(devs) * 1e3

    for tag, one_seed in (("balanced", False), ("one_seed", True)):
        progress(f"north_star {tag}: set-up")
        holders, base = multi.fanout_plan(d.Xfer, world, n_layers, size, one_seed)
        D = sum(x.size for x in base)
        my_dsts = sorted({(x.dst_node, x.layer) for x in base if x.dst_node == rank})
        fx = multi.FdExchange(rank, f"ns{tag}" + os.environ.get("MASTER_PORT", "0"))
        barrier()
        eng = d.Engine(multi.device_of_node(world, owner, rank, local_rank))
        eng.set_option("alloc_vmm", 2)               # shareable VMM layers: the same buffers serve unicast, relay and NVLS
        eng.set_option("mcast", 0)
        for l, s in holders.items():
            if s == rank:
                eng.alloc(rank, l, size)
                eng.fill(rank, l, d.FILL_PATTERN, SEED)
        for (_n, l) in my_dsts:
            eng.alloc(rank, l, size)
            eng.fill(rank, l, d.FILL_POISON)
        store = multi.SharedStore(eng, dist, world, rank, owner, fx=fx)
        r = {"deliveries": len(base), "delivered_bytes": D}
        # T*: busiest port of the placement (a layer with one holder has no sender choice); multicast/relay: egress once
        eg, ing, eg_once = [0] * world, [0] * world, [0] * world
        seen = set()
        for x in base:
            eg[x.src_node] += x.size
            ing[x.dst_node] += x.size
            if (x.src_node, x.layer) not in seen:
                seen.add((x.src_node, x.layer))
                eg_once[x.src_node] += x.size
        t_uni = max(max(eg), max(ing)) / (LINK_GBPS * 1e9) * 1e3
        t_once = max(max(eg_once), max(ing)) / (LINK_GBPS * 1e9) * 1e3
        r["t_star_unicast_ms"], r["t_star_ms"] = t_uni, t_once
        # ---- ours: dissem_plan_optimize (relay chains when they unload the busiest port, push/pull per transfer) -----
        xs, pieces = d.plan_optimize(base, gpus)
        store.share(xs)
        run = Runner(eng, multi, xs, barrier, True, owner=owner, rank=rank)
        ms, dev_ms = timed(run)
        ok = oracle_check(eng, co, o, my_dsts, size, expected) and eng.get_option("wait_timeouts") == 0
        flag = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        r.update({"path": f"relay chains, {pieces} pieces; {run.describe()}" if pieces else "unicast, push/pull per transfer",
                  "completion_ms": ms, "device_ms": dev_ms, "aggregate_GBps": D / ms / 1e6,
                  "frac_770": t_once / ms, "frac_900": t_once * LINK_GBPS / LINK_NOMINAL_GBPS / ms, "verified": bool(flag.item()),
                  "transfers_pulled": sum(1 for x in xs if x.flags & d.XFER_PULL)})
        if pieces:
            r["relay_ms"] = ms
        run.free()
        progress(f"north_star {tag}: ours {ms:.1f} ms")
        # ---- NVLS multicast on the same buffers ---------------------------------------------------------------------
        for (n, l) in my_dsts:                       # poison + fresh tables, so that the check below is not vacuous
            eng.fill(n, l, d.FILL_POISON)
            eng.reset(n, l)
        barrier()
        store.share(base)
        groups = multi.mcast_groups(base, owner)
        mc_ok, why = multi.setup_mcast(eng, fx, dist, groups, owner, rank)
        if mc_ok:
            eng.set_option("mcast", 3)
            run = Runner(eng, multi, base, barrier, True)
            ms_mc, dev_mc = timed(run)
            ok = oracle_check(eng, co, o, my_dsts, size, expected)
            gflag = torch.tensor([eng.get_option("mcast_groups")], device="cuda")
            dist.all_reduce(gflag, op=dist.ReduceOp.SUM)
            vflag = torch.tensor([1 if ok else 0], device="cuda")
            dist.all_reduce(vflag, op=dist.ReduceOp.MIN)
            r.update({"mcast_ms": ms_mc, "mcast_device_ms": dev_mc, "mcast_frac_770": t_once / ms_mc, "mcast_groups": len(groups),
                      "mcast_groups_compiled": int(gflag.item()), "mcast_verified": bool(vflag.item())})
            run.free()
            progress(f"north_star {tag}: multicast {ms_mc:.1f} ms")
        else:
            r["mcast_ms"] = None
            r["mcast_note"] = why[:200]
        barrier()
        eng.close()
        fx.close()
        barrier()                                    # every process has dropped its mappings: the layers' memory is free again
        torch.cuda.empty_cache()
        # ---- NCCL renditions of the same plan (buffers of their own; the engine's are freed) --------------------------
        plan = [(l, s, sorted(x.dst_node for x in base if x.layer == l)) for l, s in sorted(holders.items())]
        plan = [p for p in plan if p[2]]
        r.update(nccl_schedule.compare(plan, size, rank, reps=max(1, reps - 1), do_bcast=True, groups_cache=nccl_groups))
        progress(f"north_star {tag}: NCCL send/recv {r.get('nccl_sendrecv_ms')} ms, broadcast {r.get('nccl_broadcast_ms')} ms")
        out[tag] = r
    return out


def run_ours(args, rank, world, local_rank):
    import importlib

    import numpy as np
    import torch

    import dissem_b200 as d
    from oracle import c_oracle as co            # the checker: used only after the timed regions, never on the product path
    from oracle import dissem_oracle as o
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

    def progress(msg):
        if rank == 0 and args.verbose:
            print(f"[bench {time.strftime('%H:%M:%S')}] {msg}", file=sys.stderr, flush=True)

    def all_max(vals):
        t = torch.tensor(list(vals), dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return t.tolist()

    def all_ok(ok):
        if dist is None:
            return bool(ok)
        f = torch.tensor([1 if ok else 0], device="cuda")
        dist.all_reduce(f, op=dist.ReduceOp.MIN)
        return bool(f.item())

    # a hung optional leg must not cost the line: at the deadline rank 0 prints what it has and every rank leaves
    def watchdog():
        LINE.setdefault("notes", []).append(f"deadline of {args.deadline} s reached; later legs are missing")
        emit(rank)
        os._exit(0)

    dog = threading.Timer(args.deadline, watchdog)
    dog.daemon = True
    dog.start()

    w = workload_for(world)
    size = w["size"]
    text = config_text(size, w["ring"])
    owner = multi.owner_round_robin(world)
    gpu_of_node = [owner(n) for n in range(N_NODES)]

    def make_plan():
        """What a leader does with the announces in: flow.go plan -> same-GPU copies -> relay / direction choice."""
        p = d.plan(text, PLAN_MODE, time_scale=TIME_SCALE)
        xs = d.parse_plan(p)
        holders = [(nd["id"], int(l)) for nd in p["nodes"] for l in nd["layers"]]
        xs = d.plan_localize(xs, gpu_of_node, holders)
        xs, pieces = d.plan_optimize(xs, gpu_of_node)
        return p, xs, pieces

    plan, xfers, pieces = make_plan()
    D = sum(x.size for x in xfers)
    eng = d.Engine(multi.device_of_node(N_NODES, owner, rank, local_rank))
    eng.set_option("variant", args.variant)
    if args.tma_cfg is not None:
        eng.set_option("tma_cfg", args.tma_cfg)
    if args.reg_cfg is not None:
        eng.set_option("reg_cfg", args.reg_cfg)
    if args.ctas_per_sm is not None:
        eng.set_option("ctas_per_sm", args.ctas_per_sm)
    if args.unit_bytes is not None:
        eng.set_option("unit_bytes", args.unit_bytes)

    # the round-1 schedule (seeded mode-1 owner pick + direction choice), timed beside the best one at N >= 2
    xfers_m1 = []
    if world > 1 and not args.no_extras:
        xfers_m1 = d.plan_directions(d.parse_plan(d.plan(text, 1, seed=SEED, policy=0)), gpu_of_node)

    # the initial store (CreateLayers, outside the clock as in the reference) + poisoned destinations
    for nd in plan["nodes"]:
        if owner(nd["id"]) != rank:
            continue
        for l, sz in nd["layers"].items():
            eng.alloc(nd["id"], int(l), sz)
            eng.fill(nd["id"], int(l), d.FILL_PATTERN, SEED)
    for x in list(xfers) + list(xfers_m1):
        if owner(x.dst_node) == rank and not eng.has(x.dst_node, x.layer):
            eng.alloc(x.dst_node, x.layer, x.total)
            eng.fill(x.dst_node, x.layer, d.FILL_POISON)
    store = multi.SharedStore(eng, dist, world, rank, owner)
    store.share(list(xfers) + list(xfers_m1))
    run = Runner(eng, multi, xfers, barrier, world > 1)
    my_bytes = sum(x.size for x in multi.my_transfers(xfers, owner, rank))
    assert run.bytes == my_bytes, (run.bytes, my_bytes)
    my_dsts = sorted({(x.dst_node, x.layer) for x in xfers if owner(x.dst_node) == rank})
    progress(f"store ready: {len(xfers)} transfers, D = {D}")

    # first dissemination: the received counters must equal the layer sizes exactly once (they are monotonic)
    run.step()
    barrier()                                    # pushes of other ranks into this rank's buffers have landed too
    rcv_ok = all(eng.received(n, l) == size for (n, l) in my_dsts)

    # ---- device-resident throughput -------------------------------------------------------------------
    with ClockSampler(local_rank) as clocks:          # started before the warm-up so that NVML is warm; filtered below
        for _ in range(max(args.warmup, 3)):
            run.step()
        clocks.wait_first()
        barrier()
        t0 = time.perf_counter()
        times = run.pipelined(args.steps)
        barrier()
        t1 = time.perf_counter()
        wall = t1 - t0
        t_region0 = t0
        clock_window = "timed region"
        if clocks.summary(t0, t1)["samples"] < 3:
            # the timed region was too short for NVML (a few ms per query): keep the identical load running, outside the
            # timing, until a few samples exist, and say so
            t_more = time.perf_counter() + 0.25
            while time.perf_counter() < t_more:
                run.step()
            clock_window = "timed region + the same launches continued for 0.25 s (region shorter than NVML's sampling period)"
        t_clk = time.perf_counter()
    dev_ms_total = sum(t[0] for t in times)
    copy_ms_avg = sum(t[1] for t in times) / max(len(times), 1)
    wall_max, dev_ms_max = all_max([wall, dev_ms_total])
    step_s = wall_max / args.steps
    dev_step_s = dev_ms_max / 1e3 / args.steps
    progress(f"timed region: {step_s * 1e3:.4f} ms/step")

    # ---- single-shot completion time: barrier, launch, wait, barrier (cmd/main.go:173-181 times ONE dissemination) ----
    n_single = max(3, min(args.steps, 11))
    singles, single_devs = [], []
    for _ in range(n_single):
        barrier()
        ts = time.perf_counter()
        dev, _c = run.step()
        barrier()
        singles.append(time.perf_counter() - ts)
        single_devs.append(dev / 1e3)
    singles = sorted(all_max(singles))
    single_devs = sorted(all_max(single_devs))
    completion_ms = singles[len(singles) // 2] * 1e3
    completion_dev_ms = single_devs[len(single_devs) // 2] * 1e3

    # ---- correctness of what was just timed, against the CPU oracle (outside the clock) ---------------------------
    t_or = time.perf_counter()
    expected = co.fill_checksum_many(o.FILL_PATTERN, sorted({l for (_n, l) in my_dsts}), size, SEED,
                                     threads=max(1, (os.cpu_count() or 8) // world))
    ok = all_ok(oracle_check(eng, co, o, my_dsts, size, expected) and rcv_ok)
    oracle_s = time.perf_counter() - t_or
    if not ok:
        LINE.update({"metric": METRIC, "verified": False, "error": "delivered layers differ from the CPU oracle"})
        emit(rank)
        raise SystemExit("bench.py: delivered layers failed verification against the oracle")
    progress(f"verified against the oracle in {oracle_s:.1f} s")

    # ---- the round-1 schedule on the same store, for comparison (N >= 2) -------------------------------------------
    m1 = None
    if xfers_m1:
        run1 = Runner(eng, multi, xfers_m1, barrier, True)
        for _ in range(3):
            run1.step()
        barrier()
        t0 = time.perf_counter()
        run1.pipelined(args.steps)
        barrier()
        (w1,) = all_max([time.perf_counter() - t0])
        eg1, ing1, _l1 = multi.link_bytes(xfers_m1, owner, world)
        m1 = {"ms_per_step": w1 / args.steps * 1e3, "value": D / (w1 / args.steps) / 1e9,
              "busiest_port_bytes": max(max(eg1), max(ing1)),
              "what": "mode 1, seeded random owner pick + dissem_plan_directions (the schedule round 1 timed)"}
        run1.free()

    # ---- end to end through the C ABI with host buffers ---------------------------------------------------
    import ctypes
    L = d.lib()
    e2e = None
    if not args.no_e2e:
        # the bytes each local source serves live in pinned host memory; a step uploads exactly the ranges the plan reads
        ranges = {}
        for x in xfers:
            if owner(x.src_node) == rank:
                ranges.setdefault((x.src_node, x.layer), []).append((x.off, x.off + x.size))
        ranges = {k: merged_ranges(v) for k, v in ranges.items()}
        pinned = {}
        for (node, layer), rs in ranges.items():
            lo, hi = rs[0][0], rs[-1][1]
            p = ctypes.c_void_p()
            assert L.dissem_host_alloc(ctypes.byref(p), hi - lo) == 0
            pinned[(node, layer)] = (p, lo)
            assert L.dissem_layer_read(eng._ctx, node, layer, p, lo, hi - lo) == 0      # the host copy the node "loaded"
        h2d = sum(b - a for rs in ranges.values() for (a, b) in rs)
        d2h = sum(((size + d.CHUNK_BYTES - 1) // d.CHUNK_BYTES) * 16 for _ in my_dsts)
        e2e_steps = max(2, min(args.steps, 5))

        def upload():
            for (node, layer), rs in ranges.items():
                p, lo = pinned[(node, layer)]
                for (a, b) in rs:
                    eng.write(node, layer, None, off=a, ptr=p.value + (a - lo), size=b - a)

        def e2e_step():
            upload()                           # asynchronous: the planning below runs while the copy engine works
            _p, pl, _pc = make_plan()
            if dist is not None:               # a peer process's kernel cannot wait for this process's upload: meet first
                barrier()
            if dist is not None and any(d.xfer_round(x) for x in pl):      # rounds are not ordered across processes
                r_ = Runner(eng, multi, pl, barrier, True, owner=owner, rank=rank)
                r_.step()
                r_.free()
            else:
                eng.wait(eng.submit(pl))
            if dist is not None:               # pushes of other ranks into this rank's buffers have landed too
                barrier()
            return eng.checksums_many(my_dsts)

        got = e2e_step()
        e2e_ok = all_ok(all(np.array_equal(g, expected[l]) for g, (_n, l) in zip(got, my_dsts)))
        barrier()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        barrier()
        (e2e_wall,) = all_max([time.perf_counter() - t0])
        e2e_step_s = e2e_wall / e2e_steps
        # the upload alone (same ranges, waited for), to split the step into copy, device and host parts
        any_src = next(iter(ranges), None)
        barrier()
        t0 = time.perf_counter()
        for _ in range(2):
            upload()
            if any_src is not None:
                eng.received(*any_src)         # waits for the node's stream
            barrier()
        (h2d_wall,) = all_max([time.perf_counter() - t0])
        h2d_ms = h2d_wall / 2 * 1e3
        tot = torch.tensor([float(h2d), float(d2h)], dtype=torch.float64, device="cuda")
        if dist is not None:
            dist.all_reduce(tot, op=dist.ReduceOp.SUM)
        h2d_all, d2h_all = [int(v) for v in tot.tolist()]
        e2e = {"value": D / e2e_step_s / 1e9, "unit": UNIT, "h2d_bytes_per_step": h2d_all, "d2h_bytes_per_step": d2h_all,
               "ms_per_step": e2e_step_s * 1e3, "steps": e2e_steps, "h2d_ms": h2d_ms, "device_ms": completion_dev_ms,
               "host_overhead_ms": max(0.0, e2e_step_s * 1e3 - h2d_ms - completion_dev_ms),
               "host_overhead_raw_ms": e2e_step_s * 1e3 - h2d_ms - completion_dev_ms, "tables_match_oracle": e2e_ok,
               "host_overhead_what": "e2e step - upload alone (timed separately) - device time; within the noise of the upload when <= 0",
               "what": "pinned host layers -> dissem_layer_write (the byte ranges the plan reads) -> dissem_plan_json -> "
                       "dissem_plan_localize -> dissem_plan_optimize -> dissem_schedule_submit -> dissem_wait -> "
                       "dissem_layers_checksums"}
        for p, _lo in pinned.values():
            L.dissem_host_free(p)
        progress(f"e2e: {e2e['value']:.1f} GB/s")

    # ---- the line so far ---------------------------------------------------------------------------------------
    peak, peak_src = peaks()
    local = world == 1
    kernel_name = "copy_csum_tma_kernel" if args.variant == 1 else "copy_csum_kernel"
    if local:
        # algorithmic bytes of the dominant kernel: local copy = 2*S HBM bytes per delivered byte
        alg_bytes = 2 * my_bytes
        achieved = alg_bytes / (copy_ms_avg / 1e3) / 1e9 if copy_ms_avg > 0 else 0.0
        traffic = None
        tpath = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tpath):
            try:
                traffic = json.load(open(tpath)).get(kernel_name + "_dram_bytes_per_launch_" + w["key"])
            except Exception:
                traffic = None
        roof = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": kernel_name, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg_bytes, "launch_ms": copy_ms_avg}
    else:
        # NVLink roofline of the PLACEMENT under the best sender choice; what the executed plan puts on
        # its busiest port is reported beside it, not used as the bound
        t_star, best_port = best_sender_t_star(plan["nodes"], plan["assignment"], gpu_of_node, size, world)
        eg, ing, loc = multi.link_bytes(xfers, owner, world)
        busiest = max(max(e, i) for e, i in zip(eg, ing))
        roof = {"bound": "nvlink", "achieved": best_port / dev_step_s / 1e9, "peak": LINK_GBPS, "unit": "GB/s",
                "frac": t_star / dev_step_s, "traffic": None, "kernel": kernel_name,
                "peak_source": "measured peer copy per direction per GPU (B200_PROFILING.md); nominal 900",
                "algorithmic_bytes_per_launch": best_port, "launch_ms": dev_step_s * 1e3,
                "t_star_ms": t_star * 1e3, "t_star_what": "best sender choice: busiest-port bytes %d / 770 GB/s" % best_port,
                "frac_of_nominal_900": t_star * LINK_GBPS / LINK_NOMINAL_GBPS / dev_step_s,
                "frac_single_shot": t_star * 1e3 / completion_dev_ms if completion_dev_ms > 0 else None,
                "executed_plan": {"egress_bytes": eg, "ingress_bytes": ing, "local_bytes": loc, "busiest_port_bytes": busiest,
                                  "busiest_port_GBps": busiest / dev_step_s / 1e9},
                "note": "achieved = best-sender busiest-port bytes / device time of a step (max over ranks)"}
    cfg = bench_config(w, world, D)                  # identical in the reference arm, key for key
    plan_info = {"kernel_variant": "tma" if args.variant == 1 else "register", "transfers": len(xfers),
                 "transfers_pulled": sum(1 for x in xfers if x.flags & d.XFER_PULL), "relay_pieces": pieces,
                 "rounds": len(run.scheds), "units_per_step": run.units, "launches_per_step": run.launches}
    LINE.update({
        "metric": METRIC, "value": D / step_s / 1e9, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": step_s * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "u8", "data": "synthetic", "config": cfg, "plan": plan_info,
        "device_ms_per_step": dev_step_s * 1e3,
        "completion_time_ms": completion_ms, "completion_device_ms": completion_dev_ms,
        "completion_what": "median of %d single-shot disseminations: barrier, launch, wait, barrier (host clock, max over ranks); "
                           "completion_device_ms = CUDA-event time of the same, max over ranks" % n_single,
        "e2e": e2e, "gpu_launches": run.launches * args.steps,
        "clocks": dict(clocks.summary(t_region0, t_clk), window=clock_window),
        "roofline": roof, "cpu_baseline": None, "verified": ok,
        "verified_what": "fused checksum tables and spot bytes of every destination == CPU oracle (oracle.c) of the seeded fill; "
                         "received counters == layer size after the first dissemination",
    })
    if m1 is not None:
        LINE["mode1_random"] = m1

    # ---- oracle parity across NVLink (N >= 2): push, pull, multicast, relay ----------------------------------------
    run.free()
    eng.close()
    torch.cuda.empty_cache()
    parity_failed = False
    if dist is not None and not args.no_parity:
        progress("parity leg")
        try:
            par = parity_leg(d, multi, co, o, torch, dist, rank, world, local_rank, barrier)
        except Exception as e:                       # noqa: BLE001 -- the other ranks may be inside a collective: leave
            LINE["parity"] = {"error": repr(e)[:300]}
            emit(rank)
            os._exit(3)
        LINE["parity"] = par
        parity_failed = any(par.get(k) is False for k in ("push", "pull", "mcast", "relay", "relay_barrier_per_round"))

    # ---- CPU baseline on rank 0 at N=1 (bounded sample of the same workload) --------------------------------------
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        try:
            tcp = CpuTcp(w)
            secs = moved = 0.0
            t_begin = time.perf_counter()
            while tcp.passes < 400 and (tcp.passes == 0 or time.perf_counter() - t_begin < 12.0):      # about 10 s of CPU work
                s_, b_ = tcp.run_pass(verify=(tcp.passes == 0))
                secs += s_
                moved += b_
            LINE["cpu_baseline"] = {"value": moved / secs / 1e9, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port",
                                    "sample": tcp.describe() + f"; completion {secs / tcp.passes * 1e3:.1f} ms per pass"}
        except Exception as e:                       # the baseline is reported, never allowed to sink the bench line
            LINE["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": os.cpu_count() or 1, "kind": "port", "sample": f"failed: {e}"}

    # ---- the north star: Llama-3-70B-shaped assignment on 8 GPUs ---------------------------------------------------
    ns_layers, ns_size = args.north_star_layers, args.north_star_size
    if dist is not None and (world == 8 or args.north_star) and not args.no_north_star:
        try:
            LINE["north_star"] = north_star(d, multi, co, o, torch, dist, rank, world, local_rank, barrier, ns_layers, ns_size,
                                            3, progress)
        except Exception as e:                       # noqa: BLE001 -- the line is worth more than the block
            LINE["north_star"] = {"error": repr(e)[:300]}
            emit(rank)
            os._exit(0)
    dog.cancel()
    emit(rank)
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    if parity_failed:
        raise SystemExit("bench.py: a path across NVLink differs from the oracle: " + json.dumps(LINE.get("parity")))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--variant", type=int, default=1, help="0 register path, 1 TMA bulk-staged path")
    ap.add_argument("--reg-cfg", type=int, default=None)
    ap.add_argument("--tma-cfg", type=int, default=None)
    ap.add_argument("--ctas-per-sm", type=int, default=None)
    ap.add_argument("--unit-bytes", type=int, default=None)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true", help="profiling runs: skip the end-to-end leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the oracle parity leg across NVLink (N >= 2)")
    ap.add_argument("--no-extras", action="store_true", help="skip the mode-1 comparison schedule")
    ap.add_argument("--no-north-star", action="store_true")
    ap.add_argument("--north-star", action="store_true", help="run the north_star block at this N too (default: N == 8 only)")
    ap.add_argument("--north-star-layers", type=int, default=80)
    ap.add_argument("--north-star-size", type=int, default=1_711_308_800)
    ap.add_argument("--deadline", type=float, default=780.0, help="seconds after which the line is printed with what exists")
    ap.add_argument("--verbose", action="store_true")
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
