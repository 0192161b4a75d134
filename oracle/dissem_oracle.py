"""CPU ORACLE — test infrastructure, NOT product code.

A numpy / pure-Python restatement of the layer-dissemination hot path of
ynishimi/distributed-llm-dissemination.  Only `tests/`, `__graft_entry__.smoke()`
and `bench.py`'s `cpu_baseline` / `--impl reference` legs may import this file;
the product path (`distributed-llm-dissemination_b200/`) never does.

What is restated, and from where (paths relative to /root/reference):

  * config schema + dummy-layer synthesis ......... cmd/config.go:14-45, :94-171
  * mode 0 plan (leader push) ...................... distributor/node.go:326-365
  * mode 1 plan (peer retransmission) .............. distributor/node.go:554-626
  * mode 3 plan (max-flow over bisected time) ...... distributor/flow.go:55-353,
                                                     distributor/node.go:1200-1288
  * transfer semantics (bytes [Offset, Offset+DataSize) of the source layer)
                                                     distributor/transport.go:308-373 (:342)
  * completion rule ................................ distributor/node.go:435-446

PARITY STATUS:
  * pinned   — the control-flow fixtures of the reference's own tests
               (distributor/node_test.go:45-104, :163-273: identity and ring
               fixtures) and the README / conf/config.json inputs: the plans
               below reproduce exactly the transfers those fixtures imply
               (tests/test_oracle_golden.py).
  * pinned by hand-derivation only — the flow plan for conf/config.json
               (t = 60 s); the reference ships no test of flow.go.
  * PARITY UNPINNED — byte content (no reference test inspects bytes), fragment
               offsets (the reference assigns them in Go map-iteration order,
               flow.go:197-212, i.e. nondeterministically; we use ascending
               (sender, layer)), and the checksum: the reference computes NO
               checksum at all, so `checksum()` below is this repo's own
               definition and is checked only against itself (numpy vs C vs CUDA).
  * The Go toolchain is absent from the build container and the GPU box, so the
    reference itself was never executed; golang.org/x/time/rate v0.14.0 (pacing
    only, never alters bytes; transport.go:407-424) is not restated.
"""
from __future__ import annotations

import json
from dataclasses import dataclass, field
from typing import Dict, Iterable, List, Optional, Tuple

import numpy as np

# --------------------------------------------------------------------------------------
# constants shared with include/dissem.h (kept literal here on purpose: the oracle must not
# read the product's headers)
# --------------------------------------------------------------------------------------
CHUNK_BYTES = 256 * 1024          # the reference's write quantum, transport.go:408
POISON_BYTE = 0xA5
GOLDEN = 0x9E3779B97F4A7C15
MASK64 = (1 << 64) - 1
INT64_MAX = (1 << 63) - 1         # math.MaxInt64 used as "infinite" capacity, flow.go:252

# LayerLocation (node.go:187-191) / SourceType (node.go:196-200)
INMEM_LAYER, DISK_LAYER, CLIENT_LAYER = 0, 1, 2
SOURCE_CLIENT, SOURCE_DISK, SOURCE_MEM = 0, 1, 2

FILL_ZERO, FILL_PATTERN, FILL_POISON = 0, 1, 2


# --------------------------------------------------------------------------------------
# synthetic layer contents
# --------------------------------------------------------------------------------------
def _mix64(z: np.ndarray) -> np.ndarray:
    """splitmix64 finaliser on a uint64 array (wrapping arithmetic)."""
    z = z.copy()
    z ^= z >> np.uint64(30)
    z *= np.uint64(0xBF58476D1CE4E5B9)
    z ^= z >> np.uint64(27)
    z *= np.uint64(0x94D049BB133111EB)
    z ^= z >> np.uint64(31)
    return z


def fill_bytes(kind: int, layer_id: int, size: int, seed: int = 0) -> np.ndarray:
    """Bytes of a synthetic layer.

    FILL_ZERO is what the reference creates (`make([]byte, layerSize)`, cmd/config.go:141,
    :161).  FILL_PATTERN is ours: little-endian 64-bit word i of layer L is
    mix64(seed + (L+1)*GOLDEN + i); a trailing partial word is truncated.
    FILL_POISON is 0xA5 everywhere (destinations are poisoned before a transfer so that a
    missed byte cannot pass a content check)."""
    if size < 0:
        size = 0                                   # cmd/config.go:100-102
    if kind == FILL_ZERO:
        return np.zeros(size, dtype=np.uint8)
    if kind == FILL_POISON:
        return np.full(size, POISON_BYTE, dtype=np.uint8)
    if kind != FILL_PATTERN:
        raise ValueError(f"unknown fill kind {kind}")
    n_words = (size + 7) // 8
    base = (seed + (layer_id + 1) * GOLDEN) & MASK64
    with np.errstate(over="ignore"):
        idx = np.arange(n_words, dtype=np.uint64) + np.uint64(base)
        words = _mix64(idx)
    return words.astype("<u8").view(np.uint8)[:size].copy()


def checksum(data: np.ndarray, chunk_bytes: int = CHUNK_BYTES) -> np.ndarray:
    """Per-chunk checksum of one whole layer -> uint64 array [n_chunks, 2] = (A, B).

    THIS REPO'S DEFINITION (the reference has none).  The layer is read as little-endian
    uint32 words, a trailing partial word zero-extended.  For chunk c (256 KiB of the
    layer, the last one possibly short):
        A_c = sum  w_i                (mod 2^64)
        B_c = sum (i + 1) * w_i       (mod 2^64)      i = word index within the LAYER
    Both are plain sums, hence independent of reduction order and additive over disjoint
    fragments; B is position-sensitive so a fragment landing at the wrong offset shows."""
    data = np.ascontiguousarray(data, dtype=np.uint8)
    size = data.size
    n_chunks = (size + chunk_bytes - 1) // chunk_bytes
    out = np.zeros((n_chunks, 2), dtype=np.uint64)
    if size == 0:
        return out
    pad = (-size) % 4
    if pad:
        data = np.concatenate([data, np.zeros(pad, dtype=np.uint8)])
    w = data.view("<u4").astype(np.uint64)
    words_per_chunk = chunk_bytes // 4
    starts = np.arange(0, w.size, words_per_chunk)
    with np.errstate(over="ignore"):
        out[:, 0] = np.add.reduceat(w, starts)
        idx = np.arange(1, w.size + 1, dtype=np.uint64)
        out[:, 1] = np.add.reduceat(w * idx, starts)
    return out


def checksum_fragment(frag: np.ndarray, layer_off: int, layer_size: int,
                      chunk_bytes: int = CHUNK_BYTES) -> np.ndarray:
    """Contribution of bytes [layer_off, layer_off+len) to the layer's per-chunk (A, B):
    same shape as checksum() of the whole layer.  Summing (mod 2^64) the contributions of
    fragments that tile the layer gives checksum(layer) — the property the fused GPU pass
    relies on when mode 3 delivers a layer as several fragments."""
    frag = np.ascontiguousarray(frag, dtype=np.uint8)
    n_chunks = (layer_size + chunk_bytes - 1) // chunk_bytes
    out = np.zeros((n_chunks, 2), dtype=np.uint64)
    if frag.size == 0:
        return out
    pos = np.arange(layer_off, layer_off + frag.size, dtype=np.uint64)
    with np.errstate(over="ignore"):
        val = frag.astype(np.uint64) << ((pos & np.uint64(3)) * np.uint64(8))
        widx = (pos >> np.uint64(2)) + np.uint64(1)
        chunk = (pos // np.uint64(chunk_bytes)).astype(np.int64)
        np.add.at(out[:, 0], chunk, val)
        np.add.at(out[:, 1], chunk, val * widx)
    return out


def digest(csums: np.ndarray) -> int:
    """Checksum of checksums: fold a [n,2] table to one u64 (order-sensitive), used by the
    full-size property tests where comparing tables would be needlessly slow to report."""
    h = 0xCBF29CE484222325
    for a, b in np.asarray(csums, dtype=np.uint64).reshape(-1, 2).tolist():
        h = ((h ^ a) * 0x100000001B3) & MASK64
        h = ((h ^ b) * 0x100000001B3) & MASK64
    return h


# --------------------------------------------------------------------------------------
# config (cmd/config.go:14-45) — both the nested schema the code parses and the flat legacy
# schema the README documents (readme.md:17-64)
# --------------------------------------------------------------------------------------
@dataclass
class NodeConf:
    id: int
    addr: str = ""
    network_bw: int = 0
    is_leader: bool = False
    sources: Dict[int, int] = field(default_factory=dict)             # SourceType -> B/s
    # layer -> (size, source_type); what CreateLayers (cmd/config.go:94-117) would hold
    layers: Dict[int, Tuple[int, int]] = field(default_factory=dict)


@dataclass
class Config:
    nodes: List[NodeConf]
    assignment: Dict[int, List[int]]        # node -> sorted layer ids (Assignment, node.go:177)
    layer_size: int = 0
    schema: str = "nested"

    @property
    def leader(self) -> NodeConf:
        for n in self.nodes:                # GetLeaderConf, cmd/config.go:64-72 (first leader wins)
            if n.is_leader:
                return n
        raise ValueError("no leader found")

    def node(self, node_id: int) -> NodeConf:
        for n in self.nodes:
            if n.id == node_id:
                return n
        raise KeyError(node_id)

    def layer_sizes(self, any_holder: bool = False) -> Dict[int, int]:
        """layer -> size as the LEADER knows it (frleader.layers, node.go:1209, flow.go:151).
        any_holder=True is this repo's extension for placements where the leader does not hold
        every layer (the announce then carries DataSize): leader first, then ascending node id."""
        out = {l: s for l, (s, _) in self.leader.layers.items()}
        if any_holder:
            for n in sorted(self.nodes, key=lambda n: n.id):
                for l, (s, _) in n.layers.items():
                    out.setdefault(l, s)
        return out


def _ci(d: dict, key: str, default=None):
    """Go's encoding/json matches field names case-insensitively (`Id` -> `ID`)."""
    if key in d:
        return d[key]
    lk = key.lower()
    for k, v in d.items():
        if k.lower() == lk:
            return v
    return default


def parse_config(text: str) -> Config:
    raw = json.loads(text)                  # fail loudly (the reference ignores the error, config.go:59)
    layer_size = int(_ci(raw, "LayerSize", 0) or 0)
    nodes: List[NodeConf] = []
    schema = "nested"
    for nd in _ci(raw, "Nodes", []) or []:
        nc = NodeConf(id=int(_ci(nd, "ID", 0)), addr=str(_ci(nd, "Addr", "") or ""),
                      network_bw=int(_ci(nd, "NetworkBW", 0) or 0),
                      is_leader=bool(_ci(nd, "IsLeader", False)))
        for k, v in (_ci(nd, "Sources", {}) or {}).items():
            nc.sources[int(k)] = int(v)
        init = _ci(nd, "InitialLayers", {}) or {}
        # flat legacy form: {"<layer>": {}} or {"<layer>": {"LayerSize": n}}
        flat = all((not v) or (_ci(v, "LayerSize") is not None and not isinstance(_ci(v, "LayerSize"), dict))
                   for v in init.values())
        if flat and init:
            schema = "flat"
            for lk, v in sorted(init.items(), key=lambda kv: int(kv[0])):
                sz = _ci(v, "LayerSize") if v else None
                sz = layer_size if sz is None else int(sz)
                nc.layers[int(lk)] = (max(sz, 0), SOURCE_CLIENT)
        else:
            for sk in sorted(init, key=int):                       # deterministic: later source wins
                for lk, v in sorted(init[sk].items(), key=lambda kv: int(kv[0])):
                    sz = int(_ci(v, "LayerSize", 0) or 0)
                    nc.layers[int(lk)] = (max(sz, 0), int(sk))     # config.go:99-102
        nodes.append(nc)
    assignment = {int(n): sorted(int(l) for l in (ls or {}))
                  for n, ls in (_ci(raw, "Assignment", {}) or {}).items()}
    return Config(nodes=nodes, assignment=assignment, layer_size=layer_size, schema=schema)


def limit_rate(nc: NodeConf, layer: int) -> int:
    """layerSrc.Meta.LimitRate = myConf.Sources[sourceType]  (cmd/config.go:111)."""
    _, st = nc.layers[layer]
    return nc.sources.get(st, 0)


# --------------------------------------------------------------------------------------
# plans
# --------------------------------------------------------------------------------------
@dataclass(frozen=True)
class Xfer:
    """One layerMsg on the wire: bytes [off, off+size) of `layer` from src to dst
    (tempLayerInfo, transport.go:47-54: LayerSize=size, TotalSize=total, Offert=off)."""
    src: int
    dst: int
    layer: int
    off: int
    size: int
    total: int


def status_of(cfg: Config) -> Dict[int, Dict[int, int]]:
    """The leader's `status` after every node announced: node -> {layer: Location}
    (node.go:252-257 for the leader itself, node.go:303 for announces).  All synthetic layers
    are in memory when no -s path is given (cmd/main.go:94-97, config.go:108)."""
    return {n.id: {l: INMEM_LAYER for l in n.layers} for n in cfg.nodes}


def plan_mode0(cfg: Config) -> List[Xfer]:
    """LeaderNode.sendLayers, node.go:326-352: the leader sends every assigned layer the
    destination did not announce in memory."""
    st = status_of(cfg)
    leader = cfg.leader
    out = []
    for node in sorted(cfg.assignment):
        for layer in cfg.assignment[node]:
            if st.get(node, {}).get(layer) == INMEM_LAYER:           # node.go:335
                continue
            if layer not in leader.layers:                           # node.go:339-342 (warns, then hangs)
                raise ValueError(f"mode 0: leader holds no layer {layer}")
            size = leader.layers[layer][0]
            out.append(Xfer(leader.id, node, layer, 0, size, size))
    return out


def mode1_owner_candidates(cfg: Config) -> Dict[Tuple[int, int], List[int]]:
    """(dest, layer) -> sorted owners it may be served from (node.go:559-587).  The reference
    takes whichever owner Go's map iteration yields first, so every member is a valid answer."""
    st = status_of(cfg)
    owners: Dict[int, List[int]] = {}
    for node in sorted(st):
        for layer in st[node]:
            owners.setdefault(layer, []).append(node)                # lazy init as mode 2, node.go:829-834
    out = {}
    for node in sorted(cfg.assignment):
        for layer in cfg.assignment[node]:
            own = owners.get(layer, [])
            if not own:
                raise ValueError(f"mode 1: nobody holds layer {layer}")
            if node in own:                                          # node.go:578-580
                continue
            out[(node, layer)] = own
    return out


def plan_mode1(cfg: Config, seed: int = 0, policy: str = "random") -> List[Xfer]:
    """RetransmitLeaderNode.sendLayers + sendRetransmit, node.go:554-626.
    policy "random": owner = candidates[mix64(seed, dest, layer) % n] (a seeded stand-in for
    the reference's map-order pick); "balanced": least bytes sent so far, ties -> lowest id."""
    cand = mode1_owner_candidates(cfg)
    sent: Dict[int, int] = {}
    out = []
    for (node, layer), own in cand.items():
        if policy == "random":
            h = int(_mix64(np.array([(seed + (node + 1) * GOLDEN + (layer + 1) * 0xD1B54A32D192ED03) & MASK64],
                                    dtype=np.uint64))[0])
            owner = own[h % len(own)]
        elif policy == "balanced":
            owner = min(own, key=lambda o: (sent.get(o, 0), o))
        else:
            raise ValueError(policy)
        size = cfg.node(owner).layers[layer][0]
        sent[owner] = sent.get(owner, 0) + size
        out.append(Xfer(owner, node, layer, 0, size, size))
    return out


def plan_mode2_initial(cfg: Config) -> List[Xfer]:
    """PullRetransmitLeaderNode.sendLayers' initial job table, node.go:810-886 with getMinLoadedSender :948-978.
    (Which jobs are stolen later depends on measured times, node.go:980-1073; any owner is then a valid sender.)"""
    st = {n.id: dict(n.layers) for n in cfg.nodes}
    owners: Dict[int, List[int]] = {}
    for node in sorted(st):
        for layer in st[node]:
            owners.setdefault(layer, []).append(node)
    sorted_layers = sorted(owners, key=lambda l: (len(owners[l]), l))                 # rarest first, :841-851
    jobs: Dict[int, List[int]] = {}
    for node in sorted(cfg.assignment):
        for layer in cfg.assignment[node]:
            if layer not in st.get(node, {}):                                          # :858 (all layers are in memory here)
                jobs.setdefault(layer, []).append(node)
    for layer in jobs:
        if layer not in owners:
            raise ValueError(f"mode 2: nobody holds layer {layer}")
    load = {n: 0 for n in sorted(st)}
    out = []
    for layer in sorted_layers:
        for dest in jobs.get(layer, []):
            best, best_rate, min_count = 0, 0, None
            for sender in sorted(load):
                if layer not in st[sender]:
                    continue
                rate = limit_rate(cfg.node(sender), layer) or INT64_MAX                # 0 = no limit, :962-966
                cnt = load[sender]
                if rate > best_rate or (rate == best_rate and (min_count is None or cnt < min_count or
                                                               (cnt == min_count and sender < best))):
                    best, best_rate, min_count = sender, rate, cnt
            load[best] += 1
            size = cfg.node(best).layers[layer][0]
            out.append(Xfer(best, dest, layer, 0, size, size))
    return out


# ---- mode 3: flow.go ------------------------------------------------------------------
K_SOURCE, K_SENDER, K_CLIENT, K_LAYER, K_RECEIVER, K_SINK = range(6)   # flow.go:14-21


class FlowGraph:
    """flowGraph, flow.go:43-144.  Vertex numbering follows addIndex order exactly (source;
    senders ascending; per sender its source types ascending; layers ascending; receivers
    ascending; sink) because Edmonds–Karp's BFS scans the adjacency-matrix row in index order
    (flow.go:297) and the fragment sizes depend on that order."""

    def __init__(self, status_meta: Dict[int, Dict[int, Tuple[int, int]]], layer_sizes: Dict[int, int],
                 network_bw: Dict[int, int], assignment: Dict[int, List[int]], time_scale: int = 1,
                 zero_rate_unlimited: bool = False, per_dest: bool = False):
        # status_meta: node -> {layer: (LimitRate, SourceType)}
        self.status = status_meta
        self.sizes = layer_sizes
        self.bw = network_bw
        self.assignment = assignment
        self.time_scale = time_scale
        self.zero_rate_unlimited = zero_rate_unlimited
        # per_dest (this repo's extension; the reference has ONE destination per layer, node.go:1078): a layer vertex per
        # (layer, destination) pair.  Otherwise one vertex per layer, keyed (layer, None).
        self.per_dest = per_dest
        self.assignment_layers = sorted({l for ls in assignment.values() for l in ls})   # flow.go:57-65
        self.layer_vertices = sorted({(l, (n if per_dest else -1)) for n, ls in assignment.items() for l in ls})
        idx = {}

        def add(key):
            if key not in idx:
                idx[key] = len(idx)
        add((K_SOURCE,))
        for n in sorted(status_meta):                                   # flow.go:85-88
            add((K_SENDER, n))
        for n in sorted(status_meta):                                   # flow.go:91-107
            for stype in sorted({m[1] for m in status_meta[n].values()}):
                add((K_CLIENT, n, stype))
        for (l, dst) in self.layer_vertices:                            # flow.go:110-113
            add((K_LAYER, l, dst))
        for n in sorted(assignment):                                    # flow.go:116-119
            add((K_RECEIVER, n))
        add((K_SINK,))                                                  # flow.go:122
        self.idx = idx
        self.n = len(idx)
        self.adj = np.zeros((self.n, self.n), dtype=np.int64)

    def _cap(self, rate: int, t: int) -> int:
        c = rate * t // self.time_scale        # time_scale == 1 reproduces `rate*time` (flow.go:236, :250, :267)
        if c > INT64_MAX:
            raise OverflowError("capacity exceeds int64 (the Go code would wrap silently)")
        return c

    def build_edge_capacity(self, t: int) -> None:                      # flow.go:221-270
        self.adj[:] = 0
        needed = set(self.assignment_layers)
        src = self.idx[(K_SOURCE,)]
        for n in self.status:
            self.adj[src, self.idx[(K_SENDER, n)]] = self._cap(self.bw.get(n, 0), t)
        for n in sorted(self.status):
            for layer in sorted(self.status[n]):       # Go: map order; ascending makes "last write wins" deterministic
                if layer not in needed:
                    continue
                rate, stype = self.status[n][layer]
                c = self.idx[(K_CLIENT, n, stype)]
                cap = INT64_MAX if (rate == 0 and self.zero_rate_unlimited) else self._cap(rate, t)
                self.adj[self.idx[(K_SENDER, n)], c] = cap
                for (l2, dst) in self.layer_vertices:
                    if l2 == layer:
                        self.adj[c, self.idx[(K_LAYER, layer, dst)]] = INT64_MAX     # flow.go:252
        sink = self.idx[(K_SINK,)]
        for n, layers in self.assignment.items():
            r = self.idx[(K_RECEIVER, n)]
            for layer in layers:
                self.adj[self.idx[(K_LAYER, layer, n if self.per_dest else -1)], r] = self.sizes[layer]   # flow.go:261
            self.adj[r, sink] = self._cap(self.bw.get(n, 0), t)         # flow.go:267

    def bfs(self, src: int, dest: int):                                 # flow.go:283-315
        visited = np.zeros(self.n, dtype=bool)
        parent = np.zeros(self.n, dtype=np.int64)
        queue = [src]
        visited[src] = True
        head = 0
        while head < len(queue):
            u = queue[head]
            head += 1
            nxt = np.nonzero((self.adj[u] > 0) & ~visited)[0]
            for ind in nxt.tolist():
                queue.append(ind)
                visited[ind] = True
                parent[ind] = u
                if ind == dest:
                    return parent, True
        return parent, False

    def update_max_flow(self, t: int) -> int:                           # flow.go:319-353
        self.build_edge_capacity(t)
        src, sink = self.idx[(K_SOURCE,)], self.idx[(K_SINK,)]
        total = 0
        adj = self.adj
        while True:
            parent, ok = self.bfs(src, sink)
            if not ok:
                return total
            path_flow = INT64_MAX
            s = sink
            while s != src:
                path_flow = min(path_flow, int(adj[parent[s], s]))
                s = int(parent[s])
            total += path_flow
            v = sink
            while v != src:
                p = int(parent[v])
                adj[p, v] -= path_flow
                adj[v, p] += path_flow
                v = p

    def get_job_assignment(self):                                       # flow.go:146-219
        required = sum(self.sizes[l] for (l, _d) in self.layer_vertices)   # each distinct layer once, :150-152 (per pair if extended)
        t_upper = 1
        while True:
            if self.update_max_flow(t_upper) >= required:
                break
            if t_upper > INT64_MAX // 2:
                raise ValueError("tUpper not found")                    # flow.go:162-165 (logs and breaks)
            t_upper *= 2
        lo, hi, t = 1, t_upper, t_upper
        while lo <= hi:                                                 # flow.go:176-187
            m = lo + (hi - lo) // 2
            if self.update_max_flow(m) < required:
                lo = m + 1
            else:
                t = min(t, m)
                hi = m - 1
        self.update_max_flow(t)                                         # flow.go:191
        jobs = []
        offset: Dict[Tuple[int, int], int] = {}
        needed = set(self.assignment_layers)
        for sender in sorted(self.status):                              # Go: map order (:197); we fix ascending
            for layer in sorted(self.status[sender]):
                if layer not in needed:
                    continue
                _, stype = self.status[sender][layer]
                for (l2, dst) in self.layer_vertices:
                    if l2 != layer:
                        continue
                    flow = int(self.adj[self.idx[(K_LAYER, layer, dst)], self.idx[(K_CLIENT, sender, stype)]])
                    if flow > 0:                                        # reverse residual edge, :205-210
                        off = offset.get((layer, dst), 0)
                        jobs.append((sender, layer, flow, off, dst))
                        offset[(layer, dst)] = off + flow
        return t, jobs


def plan_mode3(cfg: Config, time_scale: int = 1, zero_rate_unlimited: bool = False,
               sizes_from_announces: bool = False, multi_dest: bool = False):
    """FlowRetransmitLeaderNode.assignJobs + sendLayers + dispatchJob, node.go:1200-1288.
    Returns (t, xfers, self_jobs).  `time_scale` = planner ticks per second (1 = the reference's
    integer seconds; NVLink-speed configs need 10^6 or t is always 1 and the plan degenerates).
    The reference accepts ONE destination per layer (node.go:1078, :1089-1093): the first
    assignment (ascending node id here) that names the layer."""
    leader = cfg.leader
    st_meta: Dict[int, Dict[int, Tuple[int, int]]] = {}
    for n in cfg.nodes:
        # The leader's own entry carries the full Meta (node.go:252-257); announces carry
        # Location and LimitRate only, so SourceType decodes as 0 (node.go:1397-1401).  CreateLayers
        # never sets Meta.SourceType either (config.go:104-112), so it is 0 everywhere.
        st_meta[n.id] = {l: (limit_rate(n, l), 0) for l in n.layers}
    sizes = cfg.layer_sizes(any_holder=sizes_from_announces)
    layer_dest: Dict[int, int] = {}
    for node in sorted(cfg.assignment):                                 # node.go:1087-1095
        for layer in cfg.assignment[node]:
            layer_dest.setdefault(layer, node)
    self_jobs, modified = [], {}
    for node in sorted(cfg.assignment):                                 # node.go:1206-1218
        for layer in cfg.assignment[node]:
            if layer in st_meta.get(node, {}):
                self_jobs.append((node, layer, sizes.get(layer, 0), 0))
            else:
                modified.setdefault(node, []).append(layer)
    if not modified:
        return 0, [], self_jobs                                         # node.go:1221-1224
    for ls in modified.values():
        for l in ls:
            if l not in sizes:
                raise ValueError(f"mode 3: leader does not know the size of layer {l} (flow.go:151)")
    g = FlowGraph(st_meta, sizes, {n.id: n.network_bw for n in cfg.nodes}, modified, time_scale,
                  zero_rate_unlimited, per_dest=multi_dest)
    t, jobs = g.get_job_assignment()
    xfers = [Xfer(s, (dst if multi_dest else layer_dest[l]), l, off, size, sizes[l]) for (s, l, size, off, dst) in jobs]
    return t, xfers, self_jobs


# --------------------------------------------------------------------------------------
# transfer semantics + completion
# --------------------------------------------------------------------------------------
def initial_store(cfg: Config, kind: int = FILL_PATTERN, seed: int = 0) -> Dict[Tuple[int, int], np.ndarray]:
    """(node, layer) -> bytes for every layer a node holds at start (CreateLayers, config.go:94-117).
    The same layer id carries the same bytes on every holder, as the reference's zeros do."""
    store = {}
    cache: Dict[Tuple[int, int], np.ndarray] = {}
    for n in cfg.nodes:
        for l, (size, _) in n.layers.items():
            if (l, size) not in cache:
                cache[(l, size)] = fill_bytes(kind, l, size, seed)
            store[(n.id, l)] = cache[(l, size)].copy()
    return store


def apply_transfers(store: Dict[Tuple[int, int], np.ndarray], xfers: Iterable[Xfer]) -> Dict[Tuple[int, int], np.ndarray]:
    """dst[layer][off:off+size] = src[layer][off:off+size] (transport.go:342; byte placement at
    the offset is what node.go:1545-1547's commented-out copy intends).  Destinations start
    poisoned.  Transfers are applied in order; the plans above never read a layer that an
    earlier transfer of the same plan writes."""
    for x in xfers:
        src = store[(x.src, x.layer)]
        if x.off < 0 or x.off + x.size > src.size:
            raise ValueError(f"transfer reads outside the source layer: {x}")
        dst = store.get((x.dst, x.layer))
        if dst is None:
            dst = np.full(x.total, POISON_BYTE, dtype=np.uint8)
            store[(x.dst, x.layer)] = dst
        dst[x.off:x.off + x.size] = src[x.off:x.off + x.size]
    return store


# --------------------------------------------------------------------------------------
# disk tier (`-s`): layers kept as files
# --------------------------------------------------------------------------------------
def disk_layer_path(storage_path: str, node: int, layer: int) -> str:
    """CreateDiskLayer's file name, cmd/config.go:135-139: <storage>/layers/<node id>/<layer id>.layer"""
    import os
    return os.path.join(storage_path, "layers", str(node), f"{layer}.layer")


def create_disk_layer(storage_path: str, node: int, layer: int, size: int, contents: Optional[np.ndarray] = None) -> str:
    """CreateDiskLayer, cmd/config.go:133-157: make the directory, write the file only if it does not exist yet.
    The reference writes `size` zero bytes (:141-142); `contents` lets a test put recognisable bytes there instead."""
    import os
    path = disk_layer_path(storage_path, node, layer)
    os.makedirs(os.path.dirname(path), mode=0o755, exist_ok=True)
    if not os.path.exists(path):
        data = np.zeros(max(size, 0), dtype=np.uint8) if contents is None else np.ascontiguousarray(contents, dtype=np.uint8)
        with open(path, "wb") as f:
            f.write(data.tobytes())
    return path


def read_disk_section(path: str, off: int, size: int) -> np.ndarray:
    """What sendLayerMsg puts on the wire for a DiskLayer (transport.go:356-364): the section
    [Offset, Offset+DataSize) of the file, through io.NewSectionReader + io.Copy."""
    with open(path, "rb") as f:
        f.seek(off)
        data = f.read(size)
    return np.frombuffer(data, dtype=np.uint8).copy()


# --------------------------------------------------------------------------------------
# LimitRate pacing: golang.org/x/time/rate v0.14.0 (go.mod:8; not under /root/reference) as
# writeWithLimit drives it (transport.go:407-424).  PARITY UNPINNED: no reference test exercises it.
# --------------------------------------------------------------------------------------
PACE_BURST = 256 * 1024            # BucketSize, transport.go:408


def limiter_write_times(size: int, rate: int, burst: int = PACE_BURST) -> List[Tuple[int, int, int]]:
    """[(release_ns, cumulative_start, n)] for every conn.Write of writeWithLimit(limitRate=rate, data of `size` bytes):
    n = min(len(data), burst); limiter.WaitN(ctx, n); write.  The limiter is the published token bucket:
    NewLimiter starts with `burst` tokens; reserveN advances tokens by elapsed*limit capped at burst, subtracts n, and
    a negative balance waits -tokens/limit seconds (truncated to whole ns, as time.Duration is).  Virtual time: a
    write takes no time, the next WaitN is called when the previous one returns."""
    if rate <= 0:
        raise ValueError("rate.Limit(0) never grants a second write: the reference hangs (SURVEY fact 9)")
    out = []
    tokens = float(burst)
    last = 0            # ns, lim.last
    now = 0
    pos = 0
    while pos < size:
        n = min(size - pos, burst)
        elapsed = max(0, now - last)
        tokens = min(float(burst), tokens + (elapsed / 1e9) * float(rate))      # advance()
        tokens -= float(n)
        wait = int(1e9 * (-tokens / float(rate))) if tokens < 0 else 0           # durationFromTokens
        last = now
        now = now + wait                                                         # WaitN sleeps until timeToAct
        out.append((now, pos, n))
        pos += n
    return out


def assignment_satisfied(cfg: Config, store: Dict[Tuple[int, int], np.ndarray]) -> bool:
    """assignmentSatisfied, node.go:435-446 (presence in memory; content is ours to add)."""
    return all((n, l) in store for n, ls in cfg.assignment.items() for l in ls)


def delivered_bytes(xfers: Iterable[Xfer]) -> int:
    return sum(x.size for x in xfers)


# --------------------------------------------------------------------------------------
# the synthetic configs of BASELINE.json / SURVEY §8(d), as reference-schema JSON
# --------------------------------------------------------------------------------------
S1, S2, S8B, S70B = 1_048_576, 67_108_864, 436_224_000, 1_711_308_800
NVLINK_BPS = 900_000_000_000


def _node(i, leader, layers: Dict[int, int], bw=NVLINK_BPS, rate=NVLINK_BPS):
    return {"Id": i, "Addr": f"gpu:{i}", "NetworkBW": bw, "IsLeader": leader, "Sources": {"2": rate},
            "InitialLayers": {"2": {str(l): {"LayerSize": s} for l, s in sorted(layers.items())}} if layers else {}}


def config_readme() -> str:
    """readme.md:17-64 verbatim in structure (flat schema, addresses shortened)."""
    return json.dumps({
        "Nodes": [{"Id": 0, "Addr": ":8080", "IsLeader": True, "InitialLayers": {"1": {}, "3": {}}},
                  {"Id": 1, "Addr": ":8081", "IsLeader": False, "InitialLayers": {"1": {}}},
                  {"Id": 2, "Addr": ":8082", "IsLeader": False, "InitialLayers": {}},
                  {"Id": 3, "Addr": ":8083", "IsLeader": False, "InitialLayers": {"3": {}}}],
        "Assignment": {"1": {"1": {}}, "2": {"1": {}, "3": {}}, "3": {"3": {}}},
        "LayerSize": S1})


def config_stage_ring(n_nodes=8, n_layers=32, size=S2, ring=True) -> str:
    """BASELINE config 2: leader 0 holds every layer; node n is assigned the
    stage 4n..4n+3; node n (1..N-1) initially holds the stage of its ring-next peer, mirroring
    node_test.go:63-66.  ring=False gives configs 3/5 (single seed)."""
    per = n_layers // n_nodes
    stage = lambda n: list(range(per * n, per * n + per))
    nodes = [_node(0, True, {l: size for l in range(n_layers)})]
    for n in range(1, n_nodes):
        nxt = n + 1 if n + 1 < n_nodes else 1
        held = {l: size for l in stage(nxt)} if (ring and n_nodes > 2) else {}
        nodes.append(_node(n, False, held))
    return json.dumps({"Nodes": nodes,
                       "Assignment": {str(n): {str(l): {} for l in stage(n)} for n in range(n_nodes)}})


def config_fanout(n_nodes=8, n_layers=80, size=S70B, one_seed=False) -> str:
    """BASELINE config 4: layer l -> {l, l+1, (l+2 if l%3>=1), (l+3 if l%3==2)} mod N,
    seeded on GPU (l+4) mod N (or all on GPU 0)."""
    holders: Dict[int, Dict[int, int]] = {n: {} for n in range(n_nodes)}
    assign: Dict[int, Dict[str, dict]] = {n: {} for n in range(n_nodes)}
    for l in range(n_layers):
        seed_node = 0 if one_seed else (l + 4) % n_nodes
        holders[seed_node][l] = size
        dsts = [l % n_nodes, (l + 1) % n_nodes]
        if l % 3 >= 1:
            dsts.append((l + 2) % n_nodes)
        if l % 3 == 2:
            dsts.append((l + 3) % n_nodes)
        for d in dsts:
            if d != seed_node:
                assign[d][str(l)] = {}
    # NB the leader (node 0) does not hold every layer here, so mode 0 cannot serve this placement
    # (node.go:339-342) and mode 3 needs sizes_from_announces; mode 1 is the reference mode for it.
    nodes = [_node(n, n == 0, holders[n]) for n in range(n_nodes)]
    return json.dumps({"Nodes": nodes, "Assignment": {str(n): a for n, a in assign.items() if a}})


# --------------------------------------------------------------------------------------
# the reference's own fixtures, rebuilt as configs
# --------------------------------------------------------------------------------------
def config_identity_fixture(n_receivers=4, layer_size=1) -> str:
    """node_test.go:74-104: leader 0 holds layers 1..4 in memory, receivers hold nothing,
    node k is assigned layer k; LayerSize = 1 (node_test.go:169)."""
    nodes = [_node(0, True, {l: layer_size for l in range(1, n_receivers + 1)})]
    nodes += [_node(k, False, {}) for k in range(1, n_receivers + 1)]
    return json.dumps({"Nodes": nodes,
                       "Assignment": {str(k): {str(k): {}} for k in range(1, n_receivers + 1)}})


def config_ring_fixture(n_receivers=4, layer_size=1) -> str:
    """node_test.go:45-72: as the identity fixture, but receiver index i (node i+1) already holds
    layer ((i-1+N) mod N)+1 (ring maths at :63-66)."""
    nodes = [_node(0, True, {l: layer_size for l in range(1, n_receivers + 1)})]
    for i in range(n_receivers):
        prev_layer = ((i - 1 + n_receivers) % n_receivers) + 1
        nodes.append(_node(i + 1, False, {prev_layer: layer_size}))
    return json.dumps({"Nodes": nodes,
                       "Assignment": {str(k): {str(k): {}} for k in range(1, n_receivers + 1)}})


def config_conf_json() -> str:
    """conf/config.json rebuilt from its description (nested schema): nodes 0-6 each hold layers 0-7
    under source "1" (disk, 209 715 200 B/s), node 7 holds nothing and is assigned all eight;
    NetworkBW 1 562 500 000 everywhere; Sources {"0": 16 257 500, "1": 209 715 200} on nodes 0-6.
    tests/test_oracle_golden.py checks it parses identically to the real file when that is present."""
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
